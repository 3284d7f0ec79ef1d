// Edge ingestion on the device: .npy and .csv files straight into device tensors, optionally
// fixed-point encoded on the fly, so that a multi-GB input neither needs a second host copy nor a full f64 image
// in HBM.  Counterparts in the reference: read_numpy / write_numpy (moose/src/storage/filesystem/numpy.rs:8-140,
// via the ndarray-npy crate) and read_csv / write_csv (moose/src/storage/filesystem/csv.rs:10-140), reached from
// LoadOp / SaveOp (moose/src/host/ops.rs:193-225).  The files are streamed through two page-locked staging buffers:
// read chunk k+1 from disk while chunk k is copied and converted.
#include <cerrno>
#include <cstdlib>

#include "common.h"

namespace {
constexpr int TPB = 256;
constexpr size_t STAGE_BYTES = 32u << 20;

enum SrcType { S_F8, S_F4, S_I8, S_I4, S_U8, S_U4 };

template <typename S>
__global__ void __launch_bounds__(TPB) to_f64_kernel(const S *__restrict__ src, double *__restrict__ dst, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i < n) dst[i] = (double)src[i];
}

struct NpyHeader {
  SrcType type;
  int elem;
  std::vector<int64_t> shape;
  size_t data_offset;
};

int fail_storage(mb_ctx *ctx, const std::string &m) { return mb_fail(ctx, MB_ERR_KERNEL, "storage: " + m); }

int parse_npy_header(mb_ctx *ctx, FILE *f, const std::string &path, NpyHeader *h) {
  unsigned char pre[12];
  if (fread(pre, 1, 10, f) != 10 || memcmp(pre, "\x93NUMPY", 6) != 0) return fail_storage(ctx, "failed to read numpy data file: " + path + ": bad magic");
  size_t hlen = pre[8] | (pre[9] << 8), pos = 10;
  if (pre[6] >= 2) {
    if (fread(pre + 10, 1, 2, f) != 2) return fail_storage(ctx, "failed to read numpy data file: " + path + ": short header");
    hlen |= ((size_t)pre[10] << 16) | ((size_t)pre[11] << 24);
    pos = 12;
  }
  std::string hdr(hlen, '\0');
  if (fread(&hdr[0], 1, hlen, f) != hlen) return fail_storage(ctx, "failed to read numpy data file: " + path + ": short header");
  h->data_offset = pos + hlen;
  auto value_of = [&](const char *key) -> std::string {
    size_t k = hdr.find(key);
    if (k == std::string::npos) return "";
    k = hdr.find(':', k);
    if (k == std::string::npos) return "";
    k++;
    while (k < hdr.size() && hdr[k] == ' ') k++;
    size_t e = k;
    if (hdr[k] == '(') e = hdr.find(')', k) + 1;
    else if (hdr[k] == '\'') e = hdr.find('\'', k + 1) + 1;
    else while (e < hdr.size() && hdr[e] != ',' && hdr[e] != '}') e++;
    return hdr.substr(k, e - k);
  };
  const std::string descr = value_of("'descr'"), forder = value_of("'fortran_order'"), shp = value_of("'shape'");
  if (descr.size() < 4) return fail_storage(ctx, "parsing failure from numpy data file: " + path);
  const std::string d = descr.substr(1, descr.size() - 2);
  if (d[0] == '>') return fail_storage(ctx, "big-endian numpy data is not supported: " + path);
  const std::string code = (d[0] == '<' || d[0] == '|' || d[0] == '=') ? d.substr(1) : d;
  if (code == "f8") { h->type = S_F8; h->elem = 8; }
  else if (code == "f4") { h->type = S_F4; h->elem = 4; }
  else if (code == "i8") { h->type = S_I8; h->elem = 8; }
  else if (code == "i4") { h->type = S_I4; h->elem = 4; }
  else if (code == "u8") { h->type = S_U8; h->elem = 8; }
  else if (code == "u4") { h->type = S_U4; h->elem = 4; }
  else return fail_storage(ctx, "invalid dtype for numpy storage read: " + d);
  if (forder.find("True") != std::string::npos) return fail_storage(ctx, "fortran-ordered numpy data is not supported: " + path);
  h->shape.clear();
  for (size_t i = 0; i < shp.size();) {
    if (isdigit((unsigned char)shp[i])) {
      char *end = nullptr;
      h->shape.push_back((int64_t)strtoll(shp.c_str() + i, &end, 10));
      i = (size_t)(end - shp.c_str());
    } else {
      i++;
    }
  }
  if (h->shape.size() > MB_MAX_RANK) return fail_storage(ctx, "numpy data of rank > 6: " + path);
  return MB_OK;
}

struct Stager {
  mb_ctx *ctx;
  void *host[2] = {nullptr, nullptr};
  void *dev[2] = {nullptr, nullptr};
  double *f64[2] = {nullptr, nullptr};
  cudaEvent_t done[2] = {nullptr, nullptr};
  explicit Stager(mb_ctx *c) : ctx(c) {}
  int init(bool need_f64) {
    for (int b = 0; b < 2; b++) {
      MB_CUDA(ctx, cudaMallocHost(&host[b], STAGE_BYTES));
      MB_CUDA(ctx, cudaMalloc(&dev[b], STAGE_BYTES));
      if (need_f64) MB_CUDA(ctx, cudaMalloc((void **)&f64[b], STAGE_BYTES / 4 * 8));  // 4-byte sources widen 2x
      MB_CUDA(ctx, cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming));
    }
    return MB_OK;
  }
  ~Stager() {
    cudaStreamSynchronize(ctx->stream);
    for (int b = 0; b < 2; b++) {
      if (host[b]) cudaFreeHost(host[b]);
      if (dev[b]) cudaFree(dev[b]);
      if (f64[b]) cudaFree(f64[b]);
      if (done[b]) cudaEventDestroy(done[b]);
    }
  }
};

// copies / converts `n` source elements staged in buffer b into the f64 view; returns the device pointer to f64 data
int widen(mb_ctx *ctx, Stager &st, int b, SrcType type, int64_t n, const double **out) {
  const unsigned g = (unsigned)((n + TPB - 1) / TPB);
  switch (type) {
    case S_F8: *out = (const double *)st.dev[b]; return MB_OK;
    case S_F4: to_f64_kernel<float><<<g, TPB, 0, ctx->stream>>>((const float *)st.dev[b], st.f64[b], n); break;
    case S_I8: to_f64_kernel<int64_t><<<g, TPB, 0, ctx->stream>>>((const int64_t *)st.dev[b], st.f64[b], n); break;
    case S_I4: to_f64_kernel<int32_t><<<g, TPB, 0, ctx->stream>>>((const int32_t *)st.dev[b], st.f64[b], n); break;
    case S_U8: to_f64_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)st.dev[b], st.f64[b], n); break;
    case S_U4: to_f64_kernel<uint32_t><<<g, TPB, 0, ctx->stream>>>((const uint32_t *)st.dev[b], st.f64[b], n); break;
  }
  MB_LAUNCH_CHECK(ctx);
  *out = st.f64[b];
  return MB_OK;
}

// dst: f64 tensor (ring < 0) or ring tensor encoded with 2^frac
int load_npy(mb_ctx *ctx, const char *path, int ring, uint32_t frac, mb_tensor **out) {
  FILE *f = fopen(path, "rb");
  if (!f) return fail_storage(ctx, std::string("failed to read numpy data file: ") + path + ": " + strerror(errno));
  NpyHeader h;
  int rc = parse_npy_header(ctx, f, path, &h);
  if (rc != MB_OK) { fclose(f); return rc; }
  mb_tensor *t = nullptr;
  rc = mb_tensor_new(ctx, ring < 0 ? MB_FLOAT64 : (mb_dtype)ring, (int)h.shape.size(), h.shape.data(), &t);
  if (rc != MB_OK) { fclose(f); return rc; }
  Stager st(ctx);
  rc = st.init(h.type != S_F8);
  const int64_t total = t->numel, per = (int64_t)(STAGE_BYTES / (size_t)h.elem);
  const size_t out_elem = ring < 0 ? 8 : (size_t)mb_elem_bytes((mb_dtype)ring);
  int b = 0;
  for (int64_t off = 0; off < total && rc == MB_OK; off += per, b ^= 1) {
    const int64_t n = std::min(per, total - off);
    cudaEventSynchronize(st.done[b]);  // the previous chunk staged in this buffer has been consumed
    if (fread(st.host[b], (size_t)h.elem, (size_t)n, f) != (size_t)n) { rc = fail_storage(ctx, std::string("failed to read numpy data file: ") + path + ": truncated"); break; }
    cudaError_t e = cudaMemcpyAsync(st.dev[b], st.host[b], (size_t)n * (size_t)h.elem, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) { rc = mb_fail(ctx, MB_ERR_CUDA, cudaGetErrorString(e)); break; }
    const double *src = nullptr;
    rc = widen(ctx, st, b, h.type, n, &src);
    if (rc != MB_OK) break;
    char *dst = (char *)t->dptr + (size_t)off * out_elem;
    if (ring < 0) {
      e = cudaMemcpyAsync(dst, src, (size_t)n * 8, cudaMemcpyDeviceToDevice, ctx->stream);
      if (e != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, cudaGetErrorString(e));
    } else {
      rc = k_fixed_encode_raw(ctx, src, (mb_dtype)ring, dst, n, ldexp(1.0, (int)frac));
    }
    cudaEventRecord(st.done[b], ctx->stream);
  }
  fclose(f);
  if (rc != MB_OK) { mb_tensor_free(ctx, t); return rc; }
  *out = t;
  return MB_OK;
}

}  // namespace

extern "C" int mb_load_npy(mb_ctx *ctx, const char *path, mb_tensor **out) {
  if (!ctx || !path || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_load_npy: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  return load_npy(ctx, path, -1, 0, out);
}

extern "C" int mb_load_npy_fixed(mb_ctx *ctx, const char *path, mb_dtype ring, uint32_t frac, mb_tensor **out) {
  if (!ctx || !path || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_load_npy_fixed: NULL argument");
  if (!mb_is_ring(ring) || frac >= 127) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_load_npy_fixed: bad ring / precision");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  return load_npy(ctx, path, (int)ring, frac, out);
}

extern "C" int mb_save_npy(mb_ctx *ctx, const mb_tensor *t, const char *path) {
  if (!ctx || !t || !path) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_save_npy: NULL argument");
  if (t->dtype != MB_FLOAT64 && t->dtype != MB_RING64) return mb_fail(ctx, MB_ERR_TYPE_MISMATCH, "mb_save_npy: float64 (<f8) or ring64 (<u8) tensors only");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  std::string shape = "(";
  for (int k = 0; k < t->rank; k++) shape += std::to_string(t->shape[k]) + (t->rank == 1 || k + 1 < t->rank ? "," : "") + (k + 1 < t->rank ? " " : "");
  shape += ")";
  std::string dict = std::string("{'descr': '") + (t->dtype == MB_FLOAT64 ? "<f8" : "<u8") + "', 'fortran_order': False, 'shape': " + shape + ", }";
  while ((10 + dict.size() + 1) % 64) dict.push_back(' ');
  dict.push_back('\n');
  FILE *f = fopen(path, "wb");
  if (!f) return fail_storage(ctx, std::string("failed to write moose value to file: '") + path + "': " + strerror(errno));
  const unsigned char pre[10] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0, (unsigned char)(dict.size() & 255), (unsigned char)(dict.size() >> 8)};
  bool ok = fwrite(pre, 1, 10, f) == 10 && fwrite(dict.data(), 1, dict.size(), f) == dict.size();
  void *stage = nullptr;
  if (cudaMallocHost(&stage, STAGE_BYTES) != cudaSuccess) { fclose(f); return mb_fail(ctx, MB_ERR_CUDA, "mb_save_npy: cannot allocate staging memory"); }
  for (size_t off = 0; off < t->nbytes && ok; off += STAGE_BYTES) {
    const size_t n = std::min(STAGE_BYTES, t->nbytes - off);
    ok = cudaMemcpyAsync(stage, (const char *)t->dptr + off, n, cudaMemcpyDeviceToHost, ctx->stream) == cudaSuccess &&
         cudaStreamSynchronize(ctx->stream) == cudaSuccess && fwrite(stage, 1, n, f) == n;
  }
  cudaFreeHost(stage);
  ok = fclose(f) == 0 && ok;
  return ok ? (int)MB_OK : fail_storage(ctx, std::string("failed to write moose value to file: '") + path + "'");
}

// read_csv (csv.rs:10-60): header row, every record parsed to f64, optional column selection by header name
extern "C" int mb_load_csv(mb_ctx *ctx, const char *path, const char *const *columns, int ncolumns, mb_tensor **out) {
  if (!ctx || !path || !out || (ncolumns > 0 && !columns)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_load_csv: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  FILE *f = fopen(path, "rb");
  if (!f) return fail_storage(ctx, std::string("could not open file: ") + path + ": " + strerror(errno));
  std::string text;
  char buf[1 << 16];
  size_t got;
  while ((got = fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, got);
  fclose(f);
  auto split_record = [](const std::string &line) {
    std::vector<std::string> cells;
    std::string cur;
    bool quoted = false;
    for (size_t i = 0; i < line.size(); i++) {
      const char ch = line[i];
      if (quoted) {
        if (ch == '"' && i + 1 < line.size() && line[i + 1] == '"') { cur.push_back('"'); i++; }
        else if (ch == '"') quoted = false;
        else cur.push_back(ch);
      } else if (ch == '"') quoted = true;
      else if (ch == ',') { cells.push_back(cur); cur.clear(); }
      else if (ch != '\r') cur.push_back(ch);
    }
    cells.push_back(cur);
    return cells;
  };
  std::vector<std::string> lines;
  for (size_t pos = 0; pos < text.size();) {
    size_t nl = text.find('\n', pos);
    if (nl == std::string::npos) nl = text.size();
    if (nl > pos && !(nl == pos + 1 && text[pos] == '\r')) lines.push_back(text.substr(pos, nl - pos));
    pos = nl + 1;
  }
  if (lines.empty()) return fail_storage(ctx, std::string("no columns found for file: ") + path);
  const std::vector<std::string> headers = split_record(lines[0]);
  std::vector<bool> keep(headers.size(), ncolumns == 0);
  for (int c = 0; c < ncolumns; c++)
    for (size_t k = 0; k < headers.size(); k++)
      if (headers[k] == columns[c]) keep[k] = true;
  int64_t ncols = 0;
  for (bool k : keep) ncols += k;
  std::vector<double> matrix;
  for (size_t r = 1; r < lines.size(); r++) {
    const std::vector<std::string> cells = split_record(lines[r]);
    for (size_t k = 0; k < headers.size() && k < cells.size(); k++) {
      if (!keep[k]) continue;
      char *end = nullptr;
      const double v = strtod(cells[k].c_str(), &end);
      while (end && *end == ' ') end++;
      if (end == cells[k].c_str() || (end && *end)) return fail_storage(ctx, "could not parse '" + cells[k] + "' to f64");
      matrix.push_back(v);
    }
  }
  const int64_t nrows = (int64_t)lines.size() - 1;
  if ((int64_t)matrix.size() != nrows * ncols) return fail_storage(ctx, std::string("could not convert data from: ") + path + " to matrix");
  const int64_t shape[2] = {nrows, ncols};
  return mb_tensor_from_host(ctx, MB_FLOAT64, 2, shape, matrix.data(), out);
}

// write_csv (csv.rs:62-140): header col_0..col_{n-1}, one record per row (rank <= 2)
extern "C" int mb_save_csv(mb_ctx *ctx, const mb_tensor *t, const char *path) {
  if (!ctx || !t || !path) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_save_csv: NULL argument");
  if (t->dtype != MB_FLOAT64) return fail_storage(ctx, std::string("cannot write unsupported tensor to csv file: ") + path);
  if (t->rank > 2) return fail_storage(ctx, "csv storage holds tensors of rank <= 2");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  std::vector<double> host((size_t)t->numel);
  MB_CHECK(mb_tensor_to_host(ctx, t, host.data(), t->nbytes));
  const int64_t cols = t->rank == 0 ? 1 : t->shape[t->rank - 1], rows = cols ? t->numel / cols : 0;
  FILE *f = fopen(path, "wb");
  if (!f) return fail_storage(ctx, std::string("failed to write moose value to file: '") + path + "': " + strerror(errno));
  for (int64_t c = 0; c < cols; c++) fprintf(f, "%scol_%lld", c ? "," : "", (long long)c);
  fputc('\n', f);
  for (int64_t r = 0; r < rows; r++) {
    for (int64_t c = 0; c < cols; c++) fprintf(f, "%s%.17g", c ? "," : "", host[(size_t)(r * cols + c)]);
    fputc('\n', f);
  }
  return fclose(f) == 0 ? (int)MB_OK : fail_storage(ctx, std::string("failed to write moose value to file: '") + path + "'");
}
