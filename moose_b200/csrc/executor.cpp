// C ABI: reader for Moose's textual computation format (host dialect) and a replay executor over the per-op
// kernel table of this library.
//
// Stand-in for the reference's executors on the other side of the boundary -- moose/src/execution/synchronous.rs:347-361
// (sequential walk in graph order) and asynchronous.rs:453-531 -- which are Rust and cannot be built in this image.
// Grammar restated from the nom parser in moose/src/textual/parsing.rs and the printer in moose/src/textual/mod.rs,
// one operation per line:
//     name = Kind{attr = value, ...}: (InTy, ...) -> OutTy (input, ...) @Host(role)
//     name = AddN: [HostRing128Tensor] -> HostRing128Tensor (a, b, c) @Host(role)
// Only lowered graphs are executable: after the lowering pass nothing but @Host operations on ring / bit / seed /
// key / shape values remains (moose/src/compilation/lowering.rs:4-6).  The (operator, operand types) -> kernel
// mapping is DispatchKernel::compile's (moose/src/kernels/mod.rs:34-106).
//
// Two modes.  All parties in one process: Send/Receive pairs (moose/src/networking/mod.rs:16-54) hand a value
// over through a mailbox keyed by the rendezvous key, like LocalSyncNetworking (networking/local.rs:10-53).
// One party per process: only the ops placed on `role` run, and Send/Receive go through a transport callback
// (the host binds it to NCCL send/recv between the parties' GPUs); a Receive is scheduled directly after its Send
// in the topological order every party computes identically, so matching transfers are posted in one order.
#include <chrono>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <queue>
#include <unordered_map>

#include "common.h"

namespace {

struct GOp {
  std::string name, kind;
  std::vector<std::pair<std::string, std::string>> attrs;
  std::vector<std::string> in_types;
  bool variadic = false;
  std::string out_type;
  std::vector<std::string> inputs;
  std::string plc_kind;
  std::vector<std::string> owners;
  const std::string *attr(const char *k) const {
    for (auto &a : attrs)
      if (a.first == k) return &a.second;
    return nullptr;
  }
};

struct ParseErr { std::string msg; };

std::string trim(const std::string &s) {
  size_t a = 0, b = s.size();
  while (a < b && isspace((unsigned char)s[a])) a++;
  while (b > a && isspace((unsigned char)s[b - 1])) b--;
  return s.substr(a, b - a);
}

inline bool is_open(char c) { return c == '(' || c == '[' || c == '{' || c == '<'; }
inline bool is_close(const std::string &s, size_t i) {
  const char c = s[i];
  if (c == ')' || c == ']' || c == '}') return true;
  return c == '>' && !(i > 0 && s[i - 1] == '-');  // "->" is not a bracket
}

// split at top-level separators (outside brackets and double quotes)
std::vector<std::string> split_top(const std::string &s, char sep = ',') {
  std::vector<std::string> out;
  std::string cur;
  int depth = 0;
  bool in_str = false;
  for (size_t i = 0; i < s.size(); i++) {
    const char ch = s[i];
    if (in_str) {
      cur.push_back(ch);
      if (ch == '\\' && i + 1 < s.size()) cur.push_back(s[++i]);
      else if (ch == '"') in_str = false;
    } else if (ch == '"') {
      in_str = true;
      cur.push_back(ch);
    } else if (is_open(ch)) {
      depth++;
      cur.push_back(ch);
    } else if (is_close(s, i)) {
      depth--;
      cur.push_back(ch);
    } else if (ch == sep && depth == 0) {
      out.push_back(trim(cur));
      cur.clear();
    } else {
      cur.push_back(ch);
    }
  }
  out.push_back(trim(cur));
  std::vector<std::string> kept;
  for (auto &p : out)
    if (!p.empty()) kept.push_back(p);
  return kept;
}

size_t matching_close(const std::string &s, size_t open_idx) {
  int depth = 0;
  bool in_str = false;
  for (size_t i = open_idx; i < s.size(); i++) {
    const char ch = s[i];
    if (in_str) {
      if (ch == '"' && s[i - 1] != '\\') in_str = false;
    } else if (ch == '"') {
      in_str = true;
    } else if (is_open(ch)) {
      depth++;
    } else if (is_close(s, i)) {
      if (--depth == 0) return i;
    }
  }
  throw ParseErr{"unbalanced brackets in '" + s.substr(0, 60) + "'"};
}

GOp parse_line(const std::string &raw) {
  const std::string line = trim(raw);
  GOp op;
  size_t i = 0;
  auto ident = [&](bool allow_dash) {
    const size_t b = i;
    while (i < line.size() && (isalnum((unsigned char)line[i]) || line[i] == '_' || (allow_dash && line[i] == '-'))) i++;
    return line.substr(b, i - b);
  };
  op.name = ident(true);
  while (i < line.size() && isspace((unsigned char)line[i])) i++;
  if (op.name.empty() || i >= line.size() || line[i] != '=') throw ParseErr{"cannot parse operation: '" + line.substr(0, 80) + "'"};
  i++;
  while (i < line.size() && isspace((unsigned char)line[i])) i++;
  op.kind = ident(false);
  if (op.kind.empty()) throw ParseErr{"missing operator in '" + line.substr(0, 80) + "'"};
  std::string rest = trim(line.substr(i));
  if (!rest.empty() && rest[0] == '{') {
    const size_t end = matching_close(rest, 0);
    for (auto &part : split_top(rest.substr(1, end - 1))) {
      const size_t eq = part.find('=');
      if (eq == std::string::npos) throw ParseErr{"bad attribute '" + part + "' in " + op.name};
      op.attrs.emplace_back(trim(part.substr(0, eq)), trim(part.substr(eq + 1)));
    }
    rest = trim(rest.substr(end + 1));
  }
  if (rest.empty() || rest[0] != ':') throw ParseErr{"expected ':' after operator in " + op.name};
  rest = trim(rest.substr(1));
  const size_t at = rest.rfind(" @");
  if (at == std::string::npos) throw ParseErr{"missing placement in " + op.name};
  const std::string plc = trim(rest.substr(at + 2));
  std::string body = trim(rest.substr(0, at));
  const size_t po = plc.find('(');
  if (po == std::string::npos || plc.back() != ')') throw ParseErr{"bad placement '" + plc + "' in " + op.name};
  op.plc_kind = plc.substr(0, po);
  op.owners = split_top(plc.substr(po + 1, plc.size() - po - 2));
  // signature, then the OPTIONAL argument list: a nullary operation may be written without "()"
  // (moose/src/textual/parsing.rs:176 `opt(argument_list)`); a type is alphanumeric with an optional <...>
  // (parsing.rs:497-507), so it never ends in ')'
  if (body.empty() || (body[0] != '(' && body[0] != '[')) throw ParseErr{"bad signature '" + body + "' in " + op.name};
  op.variadic = body[0] == '[';
  const size_t se = matching_close(body, 0);
  op.in_types = split_top(body.substr(1, se - 1));
  std::string after = trim(body.substr(se + 1));
  if (after.size() < 2 || after[0] != '-' || after[1] != '>') throw ParseErr{"missing '->' in signature of " + op.name};
  after = trim(after.substr(2));
  size_t te = 0;
  while (te < after.size() && isalnum((unsigned char)after[te])) te++;
  if (te == 0) throw ParseErr{"missing output type in " + op.name};
  if (te < after.size() && after[te] == '<') {
    const size_t gt = after.find('>', te);
    if (gt == std::string::npos) throw ParseErr{"unbalanced '<' in the output type of " + op.name};
    te = gt + 1;
  }
  op.out_type = after.substr(0, te);
  const std::string tail = trim(after.substr(te));
  if (!tail.empty()) {
    if (tail[0] != '(' || matching_close(tail, 0) != tail.size() - 1) throw ParseErr{"bad input list '" + tail + "' in " + op.name};
    op.inputs = split_top(tail.substr(1, tail.size() - 2));
  }
  return op;
}

// ---- values -----------------------------------------------------------------------------------------------
enum ValKind { V_NONE = -1, V_TENSOR = 0, V_BYTES = 1, V_SHAPE = 2, V_UNIT = 3, V_STRING = 4 };

struct Val {
  ValKind kind = V_NONE;
  mb_tensor *t = nullptr;
  std::vector<uint8_t> bytes;
  std::vector<int64_t> shape;
  std::string str;
};

struct ExecErr { int code; std::string msg; };
[[noreturn]] void fail(int code, const std::string &m) { throw ExecErr{code, m}; }

bool parse_u128(const std::string &s, u128 *out) {
  std::string t = trim(s);
  bool negv = false;
  size_t i = 0;
  if (!t.empty() && (t[0] == '-' || t[0] == '+')) { negv = t[0] == '-'; i = 1; }
  if (i >= t.size()) return false;
  u128 v = 0;
  for (; i < t.size(); i++) {
    if (t[i] == '_') continue;
    if (!isdigit((unsigned char)t[i])) return false;
    v = v * 10 + (u128)(t[i] - '0');
  }
  *out = negv ? (u128)0 - v : v;
  return true;
}

int64_t to_i64(const GOp &op, const char *key) {
  const std::string *v = op.attr(key);
  if (!v) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": missing attribute " + key);
  char *end = nullptr;
  const long long r = strtoll(v->c_str(), &end, 10);
  if (end == v->c_str()) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": attribute " + key + " is not an integer");
  return (int64_t)r;
}

std::string unquote(const std::string &v) {
  if (v.size() >= 2 && v.front() == '"' && v.back() == '"') {
    std::string o;
    for (size_t i = 1; i + 1 < v.size(); i++) {
      if (v[i] == '\\' && i + 2 < v.size()) i++;
      o.push_back(v[i]);
    }
    return o;
  }
  return v;
}

// sync_key / rendezvous_key: 32 hex digits, or a byte list zero-padded to 16 (SyncKey::try_from, host/prim.rs:98-111)
void hex16(const GOp &op, const char *key, uint8_t out[16]) {
  const std::string *pv = op.attr(key);
  if (!pv) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": missing attribute " + key);
  const std::string v = trim(*pv);
  memset(out, 0, 16);
  if (!v.empty() && v[0] == '[') {
    int n = 0;
    for (auto &p : split_top(v.substr(1, v.size() - 2))) {
      if (n >= 16) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": key longer than 16 bytes");
      out[n++] = (uint8_t)strtoul(p.c_str(), nullptr, 10);
    }
    return;
  }
  if (v.size() > 32 || v.size() % 2) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": bad hex key");
  for (size_t i = 0; i < v.size(); i += 2) out[i / 2] = (uint8_t)strtoul(v.substr(i, 2).c_str(), nullptr, 16);
}

// nested list literal -> flat scalars + shape
void parse_nested(const std::string &s, std::vector<std::string> &flat, std::vector<int64_t> &shape, int depth = 0) {
  const std::string t = trim(s);
  if (!t.empty() && t[0] == '[') {
    const auto parts = split_top(t.substr(1, matching_close(t, 0) - 1));
    if ((int)shape.size() <= depth) shape.push_back((int64_t)parts.size());
    else if (shape[depth] != (int64_t)parts.size()) fail(MB_ERR_INVALID_ARGUMENT, "ragged tensor literal");
    for (auto &p : parts) parse_nested(p, flat, shape, depth + 1);
  } else {
    flat.push_back(t);
  }
}

bool split_constant(const std::string &value, std::string *ty, std::string *payload) {
  const std::string v = trim(value);
  const size_t po = v.find('(');
  if (po == std::string::npos || v.back() != ')') return false;
  *ty = v.substr(0, po);
  *payload = trim(v.substr(po + 1, v.size() - po - 2));
  return true;
}

int dtype_of(const std::string &ty) {
  if (ty == "HostRing64Tensor" || ty == "Ring64Tensor") return MB_RING64;
  if (ty == "HostRing128Tensor" || ty == "Ring128Tensor") return MB_RING128;
  if (ty == "HostBitTensor" || ty == "BitTensor") return MB_BIT;
  if (ty == "HostFloat64Tensor" || ty == "Float64Tensor") return MB_FLOAT64;
  return -1;
}

}  // namespace

struct mb_graph {
  std::vector<GOp> ops;
  std::vector<int> order;  // topological order, every Receive right after its Send
  std::unordered_map<std::string, int> index;
};

struct mb_exec {
  mb_ctx *ctx;
  const mb_graph *g;
  bool has_role;
  std::string role;
  mb_transport_fn transport;
  void *transport_user;
  bool stream_ordered;  // transport capability bit 0 (mode 5): device payloads are ordered on the context's stream
  std::vector<Val> values;
  std::unordered_map<std::string, Val> bound;    // overrides / Input arguments by op name resp. arg_name
  std::map<std::string, int> outputs, saved;     // tag / key -> op index holding the value
  std::unordered_map<std::string, int> mailbox;  // rendezvous key -> index of the sent value
  uint64_t executed, bytes_sent;
  double transport_seconds, sync_seconds;  // host wall time inside the transport / waiting for the stream around it
  bool keep_values;                        // false: a value is released after its last consumer ran
  std::deque<mb_tensor *> in_flight;       // device payloads handed to the transport, oldest first (each holds a reference)
  uint64_t sends_reported;                 // completed device sends the transport has reported so far
};

namespace {

void release(mb_ctx *ctx, Val &v) {
  if (v.t) mb_tensor_free(ctx, v.t);
  v = Val();
}

void toposort(mb_graph *g) {
  const int n = (int)g->ops.size();
  std::unordered_map<std::string, int> send_of;
  for (int i = 0; i < n; i++)
    if (g->ops[i].kind == "Send") {
      const std::string *k = g->ops[i].attr("rendezvous_key");
      if (k) send_of[*k] = i;
    }
  std::vector<std::vector<int>> users((size_t)n);
  std::vector<int> indeg((size_t)n, 0);
  for (int i = 0; i < n; i++) {
    std::vector<int> deps;
    for (auto &in : g->ops[i].inputs) {
      auto it = g->index.find(in);
      if (it != g->index.end()) deps.push_back(it->second);
    }
    if (g->ops[i].kind == "Receive") {
      const std::string *k = g->ops[i].attr("rendezvous_key");
      if (k && send_of.count(*k)) deps.push_back(send_of[*k]);
    }
    indeg[(size_t)i] = (int)deps.size();
    for (int d : deps) users[(size_t)d].push_back(i);
  }
  typedef std::pair<int, int> PI;  // (priority, op)
  std::priority_queue<PI, std::vector<PI>, std::greater<PI>> ready;
  for (int i = 0; i < n; i++)
    if (indeg[(size_t)i] == 0) ready.push(PI(i, i));
  while (!ready.empty()) {
    const int i = ready.top().second;
    ready.pop();
    g->order.push_back(i);
    for (int u : users[(size_t)i])
      if (--indeg[(size_t)u] == 0) {
        const bool pair = g->ops[(size_t)i].kind == "Send" && g->ops[(size_t)u].kind == "Receive";
        ready.push(PI(pair ? -1 : u, u));
      }
  }
  if ((int)g->order.size() != n) throw ParseErr{"computation has a cycle or a dangling input"};
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct Runner {
  mb_exec *e;
  mb_ctx *c;
  int xfer(int is_send, const std::string &peer, void *buf, size_t n, int dev) {
    const double t0 = now_s();
    const int rc = e->transport(e->transport_user, is_send, peer.c_str(), buf, n, dev);
    e->transport_seconds += now_s() - t0;
    return rc;
  }
  void sync() {
    if (e->stream_ordered) return;
    const double t0 = now_s();
    ck(mb_ctx_sync(c));
    e->sync_seconds += now_s() - t0;
  }
  explicit Runner(mb_exec *e_) : e(e_), c(e_->ctx) {}

  void ck(int rc) {
    if (rc != MB_OK) throw ExecErr{rc, mb_last_error(c)};
  }
  const Val &in(const GOp &op, size_t k) {
    if (k >= op.inputs.size()) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": missing operand");
    auto it = e->g->index.find(op.inputs[k]);
    if (it == e->g->index.end() || e->values[(size_t)it->second].kind == V_NONE)
      fail(MB_ERR_KERNEL, op.name + ": operand " + op.inputs[k] + " has no value");
    return e->values[(size_t)it->second];
  }
  const mb_tensor *tin(const GOp &op, size_t k) {
    const Val &v = in(op, k);
    if (v.kind != V_TENSOR) fail(MB_ERR_TYPE_MISMATCH, op.name + ": operand " + op.inputs[k] + " is not a tensor");
    return v.t;
  }
  const std::vector<int64_t> &shin(const GOp &op, size_t k) {
    const Val &v = in(op, k);
    if (v.kind != V_SHAPE) fail(MB_ERR_TYPE_MISMATCH, op.name + ": operand " + op.inputs[k] + " is not a shape");
    return v.shape;
  }
  const uint8_t *bin(const GOp &op, size_t k) {
    const Val &v = in(op, k);
    if (v.kind != V_BYTES || v.bytes.size() != 16) fail(MB_ERR_TYPE_MISMATCH, op.name + ": operand " + op.inputs[k] + " is not a key / seed");
    return v.bytes.data();
  }
  int out_dtype(const GOp &op) {
    const int d = dtype_of(op.out_type);
    if (d < 0 || d == MB_FLOAT64) fail(MB_ERR_TYPE_MISMATCH, op.name + ": no device dtype for type " + op.out_type);
    return d;
  }
  Val tval(mb_tensor *t) {
    Val v;
    v.kind = V_TENSOR;
    v.t = t;
    return v;
  }
  Val copy(const Val &s) {
    Val v = s;
    if (v.t) mb_tensor_retain(v.t);
    return v;
  }
  int64_t max_value(const GOp &op) { return op.attr("max_value") ? to_i64(op, "max_value") : -1; }

  Val constant(const GOp &op, const std::string &literal) {
    std::string ty, payload;
    if (!split_constant(literal, &ty, &payload)) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": bad constant");
    Val v;
    if (ty == "HostPrfKey" || ty == "HostSeed" || ty == "PrfKey" || ty == "Seed") {
      v.kind = V_BYTES;
      for (size_t i = 0; i + 1 < payload.size(); i += 2) v.bytes.push_back((uint8_t)strtoul(payload.substr(i, 2).c_str(), nullptr, 16));
      v.bytes.resize(16, 0);
      return v;
    }
    if (ty == "HostString" || ty == "String") {
      v.kind = V_STRING;
      v.str = unquote(payload);
      return v;
    }
    if (ty == "HostShape" || ty == "Shape") {
      v.kind = V_SHAPE;
      const std::string t = trim(payload);
      for (auto &p : split_top(t.substr(1, t.size() - 2))) v.shape.push_back((int64_t)strtoll(p.c_str(), nullptr, 10));
      return v;
    }
    const int dt = dtype_of(ty);
    if (dt < 0) fail(MB_ERR_UNIMPLEMENTED, op.name + ": constants of type " + ty + " are not part of the host-dialect subset");
    std::vector<std::string> flat;
    std::vector<int64_t> shape;
    parse_nested(payload, flat, shape);
    if (shape.size() > MB_MAX_RANK) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": constant rank too large");
    mb_tensor *t = nullptr;
    if (dt == MB_FLOAT64) {
      std::vector<double> d;
      for (auto &s : flat) d.push_back(strtod(s.c_str(), nullptr));
      ck(mb_tensor_from_host(c, MB_FLOAT64, (int)shape.size(), shape.data(), d.data(), &t));
    } else if (dt == MB_BIT) {
      std::vector<uint8_t> packed((flat.size() + 7) / 8 + 1, 0);
      for (size_t i = 0; i < flat.size(); i++)
        if (strtoul(flat[i].c_str(), nullptr, 10) & 1) packed[i >> 3] |= (uint8_t)(1u << (i & 7));
      ck(mb_tensor_from_host(c, MB_BIT, (int)shape.size(), shape.data(), packed.data(), &t));
    } else {
      std::vector<uint64_t> w;
      for (auto &s : flat) {
        u128 x;
        if (!parse_u128(s, &x)) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": bad ring literal " + s);
        w.push_back((uint64_t)x);
        if (dt == MB_RING128) w.push_back((uint64_t)(x >> 64));
      }
      ck(mb_tensor_from_host(c, (mb_dtype)dt, (int)shape.size(), shape.data(), w.data(), &t));
    }
    return tval(t);
  }

  // the transport: a 10-word header (kind, dtype, rank, shape[6], nbytes) then the payload
  void send(const Val &v, const std::string &peer) {
    int64_t hdr[10] = {0};
    hdr[0] = v.kind;
    const void *payload = nullptr;
    size_t nbytes = 0;
    int on_device = 0;
    if (v.kind == V_TENSOR) {
      hdr[1] = mb_tensor_dtype(v.t);
      hdr[2] = mb_tensor_rank(v.t);
      for (int k = 0; k < hdr[2]; k++) hdr[3 + k] = mb_tensor_shape(v.t)[k];
      nbytes = mb_tensor_nbytes(v.t);
      payload = mb_tensor_device_ptr(v.t);
      on_device = 1;
      sync();  // the payload is read by another stream / library
    } else if (v.kind == V_BYTES) {
      nbytes = v.bytes.size();
      payload = v.bytes.data();
    } else if (v.kind == V_STRING) {
      nbytes = v.str.size();
      payload = v.str.data();
    } else if (v.kind == V_SHAPE) {
      if (v.shape.size() > 6) fail(MB_ERR_INVALID_ARGUMENT, "cannot send a shape of rank > 6");
      hdr[2] = (int64_t)v.shape.size();
      for (size_t k = 0; k < v.shape.size(); k++) hdr[3 + k] = v.shape[k];
    } else if (v.kind != V_UNIT) {
      fail(MB_ERR_KERNEL, "cannot send an empty value");
    }
    hdr[9] = (int64_t)nbytes;
    if (xfer(1, peer, hdr, sizeof(hdr), 0) != 0) fail(MB_ERR_KERNEL, "transport: send of header to " + peer + " failed");
    if (nbytes && xfer(1, peer, const_cast<void *>(payload), nbytes, on_device) != 0)
      fail(MB_ERR_KERNEL, "transport: send of payload to " + peer + " failed");
    if (nbytes && on_device) e->in_flight.push_back(mb_tensor_retain(v.t));  // the transport may still be reading it
    e->bytes_sent += nbytes;
    reap(false);
  }
  // Device payloads stay referenced until the transport reports the send complete (mode 2: number of completed
  // device sends so far; mode 3: wait for all of them).
  void reap(bool wait_all) {
    if (e->in_flight.empty() && !wait_all) return;
    int64_t done = 0;
    if (xfer(wait_all ? 3 : 2, "", &done, sizeof(done), 0) != 0) fail(MB_ERR_KERNEL, "transport: completion query failed");
    while (e->sends_reported < (uint64_t)done && !e->in_flight.empty()) {
      mb_tensor_free(c, e->in_flight.front());
      e->in_flight.pop_front();
      e->sends_reported++;
    }
  }
  Val recv(const std::string &peer) {
    int64_t hdr[10] = {0};
    if (xfer(0, peer, hdr, sizeof(hdr), 0) != 0) fail(MB_ERR_KERNEL, "transport: receive of header from " + peer + " failed");
    Val v;
    v.kind = (ValKind)hdr[0];
    if (hdr[9] < 0) fail(MB_ERR_KERNEL, "transport: negative payload size in a header from " + peer);
    const size_t nbytes = (size_t)hdr[9];
    if ((v.kind == V_BYTES || v.kind == V_STRING) && nbytes > 4096)   // host payloads travel in 4096-byte transport slots
      fail(MB_ERR_KERNEL, "transport: host payload larger than a transport slot in a header from " + peer);
    if (v.kind == V_TENSOR) {
      if (hdr[2] < 0 || hdr[2] > MB_MAX_RANK) fail(MB_ERR_KERNEL, "transport: bad tensor header");
      if (hdr[1] != MB_RING64 && hdr[1] != MB_RING128 && hdr[1] != MB_BIT && hdr[1] != MB_FLOAT64)
        fail(MB_ERR_KERNEL, "transport: unknown dtype in a tensor header from " + peer);
      // receive buffer outside the stream order: the transfer can start while this party's kernels still run
      ck(e->stream_ordered ? mb_tensor_new(c, (mb_dtype)hdr[1], (int)hdr[2], hdr + 3, &v.t)
                           : mb_tensor_new_detached(c, (mb_dtype)hdr[1], (int)hdr[2], hdr + 3, &v.t));
      if (mb_tensor_nbytes(v.t) != nbytes) fail(MB_ERR_KERNEL, "transport: payload size does not match the tensor header");
      if (nbytes && xfer(0, peer, mb_tensor_device_ptr(v.t), nbytes, 1) != 0)
        fail(MB_ERR_KERNEL, "transport: receive of payload from " + peer + " failed");
    } else if (v.kind == V_BYTES || v.kind == V_STRING) {
      std::vector<uint8_t> buf(nbytes + 1, 0);
      if (nbytes && xfer(0, peer, buf.data(), nbytes, 0) != 0)
        fail(MB_ERR_KERNEL, "transport: receive of payload from " + peer + " failed");
      if (v.kind == V_BYTES) v.bytes.assign(buf.begin(), buf.begin() + (long)nbytes);
      else v.str.assign((const char *)buf.data(), nbytes);
    } else if (v.kind == V_SHAPE) {
      if (hdr[2] < 0 || hdr[2] > 6) fail(MB_ERR_KERNEL, "transport: bad shape header");
      v.shape.assign(hdr + 3, hdr + 3 + hdr[2]);
    } else if (v.kind != V_UNIT) {
      fail(MB_ERR_KERNEL, "transport: unknown value kind");
    }
    return v;
  }

  Val execute(const GOp &op, int idx) {
    const std::string &k = op.kind;
    mb_tensor *t = nullptr;
    if (k == "PrfKeyGen") {
      Val v;
      v.kind = V_BYTES;
      v.bytes.resize(16);
      ck(mb_prf_key_gen(c, v.bytes.data()));
      return v;
    }
    if (k == "DeriveSeed") {
      uint8_t sync[16];
      hex16(op, "sync_key", sync);
      Val v;
      v.kind = V_BYTES;
      v.bytes.resize(16);
      ck(mb_derive_seed(c, sync, bin(op, 0), v.bytes.data()));
      return v;
    }
    if (k == "Constant") {
      const std::string *lit = op.attr("value");
      if (!lit) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": Constant without value");
      return constant(op, *lit);
    }
    if (k == "Input") {
      const std::string *an = op.attr("arg_name");
      auto it = e->bound.find(an ? unquote(*an) : op.name);
      if (it == e->bound.end()) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": no argument bound for this Input");
      return copy(it->second);
    }
    if (k == "Identity") return copy(in(op, 0));
    if (k == "RingFixedpointEncode" || k == "RingFixedpointDecode") {
      if (to_i64(op, "scaling_base") != 2) fail(MB_ERR_UNIMPLEMENTED, op.name + ": only scaling_base = 2");
      if (k == "RingFixedpointEncode") ck(mb_fixedpoint_encode(c, (mb_dtype)out_dtype(op), (uint32_t)to_i64(op, "scaling_exp"), tin(op, 0), &t));
      else ck(mb_fixedpoint_decode(c, (uint32_t)to_i64(op, "scaling_exp"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "Shape") {
      const mb_tensor *x = tin(op, 0);
      Val v;
      v.kind = V_SHAPE;
      v.shape.assign(mb_tensor_shape(x), mb_tensor_shape(x) + mb_tensor_rank(x));
      return v;
    }
    if (k == "Fill") {
      const std::string *lit = op.attr("value");
      std::string ty, payload;
      u128 val = 0;
      if (!lit || !split_constant(*lit, &ty, &payload) || !parse_u128(payload, &val)) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": bad fill value");
      const auto &sh = shin(op, 0);
      ck(mb_fill(c, (mb_dtype)out_dtype(op), (int)sh.size(), sh.data(), (uint64_t)val, (uint64_t)(val >> 64), &t));
      return tval(t);
    }
    if (k == "SampleSeeded") {
      const auto &sh = shin(op, 0);
      ck(mb_sample_seeded(c, (mb_dtype)out_dtype(op), (int)sh.size(), sh.data(), bin(op, 1), max_value(op), &t));
      return tval(t);
    }
    if (k == "Sample") {
      const auto &sh = shin(op, 0);
      ck(mb_sample(c, (mb_dtype)out_dtype(op), (int)sh.size(), sh.data(), max_value(op), &t));
      return tval(t);
    }
    typedef int (*Bin)(mb_ctx *, const mb_tensor *, const mb_tensor *, mb_tensor **);
    static const std::pair<const char *, Bin> bins[] = {{"Add", mb_add}, {"Sub", mb_sub}, {"Mul", mb_mul}, {"Dot", mb_dot}, {"Div", mb_div},
                                                        {"Xor", mb_xor}, {"And", mb_and}, {"Or", mb_or}, {"Less", mb_less}, {"Greater", mb_greater}};
    for (auto &b : bins)
      if (k == b.first) {
        ck(b.second(c, tin(op, 0), tin(op, 1), &t));
        return tval(t);
      }
    typedef int (*Un)(mb_ctx *, const mb_tensor *, mb_tensor **);
    static const std::pair<const char *, Un> uns[] = {{"Neg", mb_neg}, {"Transpose", mb_transpose}, {"Diag", mb_diag}, {"Sign", mb_sign},
                                                      {"Cast", mb_cast_ring128_to_ring64}};
    for (auto &u : uns)
      if (k == u.first) {
        ck(u.second(c, tin(op, 0), &t));
        return tval(t);
      }
    if (k == "Shl" || k == "Shr") {
      ck((k == "Shl" ? mb_shl : mb_shr)(c, (uint64_t)to_i64(op, "amount"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "AddN" || k == "Maximum" || k == "Concat") {
      std::vector<const mb_tensor *> xs;
      for (size_t i = 0; i < op.inputs.size(); i++) xs.push_back(tin(op, i));
      if (k == "AddN") ck(mb_add_n(c, xs.data(), (int)xs.size(), &t));
      else if (k == "Maximum") ck(mb_maximum(c, xs.data(), (int)xs.size(), &t));
      else ck(mb_concat(c, (int)to_i64(op, "axis"), xs.data(), (int)xs.size(), &t));
      return tval(t);
    }
    if (k == "Sum") {
      const std::string *a = op.attr("axis");
      ck(mb_sum(c, tin(op, 0), (!a || *a == "None") ? -1 : (int)to_i64(op, "axis"), &t));
      return tval(t);
    }
    if (k == "BitDecompose") {
      ck((dtype_of(op.out_type) == MB_BIT ? mb_bit_decompose : mb_bit_decompose_ring)(c, tin(op, 0), &t));
      return tval(t);
    }
    if (k == "BitExtract") {
      ck(mb_bit_extract(c, (uint64_t)to_i64(op, "bit_idx"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "RingInject") {
      ck(mb_ring_inject(c, (mb_dtype)out_dtype(op), (uint64_t)to_i64(op, "bit_idx"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "ShlDim") {
      ck(mb_shl_dim(c, (uint64_t)to_i64(op, "amount"), (uint64_t)to_i64(op, "bit_length"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "IndexAxis") {
      ck(mb_index_axis(c, (int)to_i64(op, "axis"), to_i64(op, "index"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "ExpandDims") {
      const std::string *a = op.attr("axis");
      if (!a) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": ExpandDims without axis");
      std::vector<int64_t> axes;
      const std::string v = trim(*a);
      for (auto &p : split_top(v.size() >= 2 && v[0] == '[' ? v.substr(1, v.size() - 2) : v)) axes.push_back((int64_t)strtoll(p.c_str(), nullptr, 10));
      ck(mb_expand_dims(c, axes.data(), (int)axes.size(), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "Squeeze") {
      ck(mb_squeeze(c, op.attr("axis") ? (int)to_i64(op, "axis") : -1, tin(op, 0), &t));
      return tval(t);
    }
    if (k == "Slice") {
      // the textual form carries ONE SliceInfoElem (slice_info_literal, parsing.rs:902-913)
      const std::string *lit = op.attr("slice");
      if (!lit || lit->size() < 2 || (*lit)[0] != '{') fail(MB_ERR_INVALID_ARGUMENT, op.name + ": bad slice literal");
      int64_t start = 0, end = 0, step = 1;
      uint8_t has_end = 0;
      bool has_start = false;
      for (auto &p : split_top(lit->substr(1, lit->size() - 2))) {
        const size_t eq = p.find('=');
        if (eq == std::string::npos) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": bad slice literal");
        const std::string f = trim(p.substr(0, eq));
        const int64_t val = (int64_t)strtoll(trim(p.substr(eq + 1)).c_str(), nullptr, 10);
        if (f == "start") { start = val; has_start = true; }
        else if (f == "end") { end = val; has_end = 1; }
        else if (f == "step") step = val;
      }
      if (!has_start) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": slice literal without start");
      const Val &x = in(op, 0);
      if (x.kind == V_SHAPE) {  // SliceOp::shape_kernel (host/ops.rs:537-546)
        if (!has_end || start < 0 || end < start || end > (int64_t)x.shape.size()) fail(MB_ERR_KERNEL, op.name + ": shape slice out of range");
        Val v;
        v.kind = V_SHAPE;
        v.shape.assign(x.shape.begin() + start, x.shape.begin() + end);
        return v;
      }
      ck(mb_slice(c, 1, &start, &end, &has_end, &step, tin(op, 0), &t));
      return tval(t);
    }
    if (k == "Select") {
      ck(mb_select(c, (int)to_i64(op, "axis"), tin(op, 1), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "Mux") {
      ck(mb_mux(c, tin(op, 0), tin(op, 1), tin(op, 2), &t));
      return tval(t);
    }
    if (k == "RingFixedpointMean") {
      ck(mb_fixedpoint_mean(c, op.attr("axis") ? (int)to_i64(op, "axis") : -1, (uint64_t)to_i64(op, "scaling_base"),
                            (uint32_t)to_i64(op, "scaling_exp"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "RingFixedpointArgmax") {
      ck(mb_fixedpoint_argmax(c, (int)to_i64(op, "axis"), tin(op, 0), &t));
      return tval(t);
    }
    if (k == "Reshape") {
      const auto &sh = shin(op, 1);
      ck(mb_reshape(c, tin(op, 0), (int)sh.size(), sh.data(), &t));
      return tval(t);
    }
    if (k == "Broadcast") {
      const auto &sh = shin(op, 0);
      ck(mb_broadcast(c, tin(op, 1), (int)sh.size(), sh.data(), &t));
      return tval(t);
    }
    if (k == "Output") {
      const std::string *tag = op.attr("tag");
      e->outputs[tag ? unquote(*tag) : op.name] = idx;
      return copy(in(op, 0));
    }
    if (k == "Load") {
      // LoadOp (host/ops.rs:193-208): key (and a column query for csv); bound values stand for in-memory storage
      const Val &key = in(op, 0);
      if (key.kind != V_STRING) fail(MB_ERR_TYPE_MISMATCH, op.name + ": Load key is not a string");
      auto it = e->bound.find(key.str);
      if (it != e->bound.end()) return copy(it->second);
      const std::string &p = key.str;
      if (dtype_of(op.out_type) != MB_FLOAT64) fail(MB_ERR_UNIMPLEMENTED, op.name + ": file storage yields HostFloat64Tensor");
      if (p.size() > 4 && p.compare(p.size() - 4, 4, ".npy") == 0) ck(mb_load_npy(c, p.c_str(), &t));
      else if (p.size() > 4 && p.compare(p.size() - 4, 4, ".csv") == 0) ck(mb_load_csv(c, p.c_str(), nullptr, 0, &t));
      else fail(MB_ERR_KERNEL, "key must provide an extension of either '.csv' or '.npy', got: " + p);
      return tval(t);
    }
    if (k == "Save") {
      const Val &key = in(op, 0);
      if (key.kind != V_STRING) fail(MB_ERR_TYPE_MISMATCH, op.name + ": Save key is not a string");
      e->saved[key.str] = idx;
      return copy(in(op, 1));
    }
    fail(MB_ERR_UNIMPLEMENTED, "operator " + k + " is not part of the host-dialect subset built here");
  }

  void run() {
    const mb_graph *g = e->g;
    const bool trace = getenv("MOOSE_B200_EXEC_TRACE") != nullptr;  // per-op host wall time (stream drained per op)
    double t_prev = now_s();
    // consumers left per value (only those that run in this process); Output / Save keep their own reference
    std::vector<int> uses(g->ops.size(), 0);
    auto is_mine = [&](const GOp &o) { return !e->has_role || (!o.owners.empty() && o.owners[0] == e->role); };
    if (!e->keep_values)
      for (auto &o : g->ops)
        if (is_mine(o))
          for (auto &in_name : o.inputs) {
            auto it = g->index.find(in_name);
            if (it != g->index.end()) uses[(size_t)it->second]++;
          }
    auto consumed = [&](const GOp &o) {
      if (e->keep_values) return;
      for (auto &in_name : o.inputs) {
        auto it = g->index.find(in_name);
        if (it == g->index.end()) continue;
        bool in_mailbox = false;
        for (auto &m : e->mailbox) in_mailbox |= m.second == it->second;
        if (--uses[(size_t)it->second] == 0 && !in_mailbox) release(c, e->values[(size_t)it->second]);
      }
    };
    for (int idx : g->order) {
      if (trace) {
        cudaStreamSynchronize(c->stream);
        const double t = now_s();
        if (idx != g->order.front() && t - t_prev > 1e-3)
          fprintf(stderr, "[exec %s] %8.3f ms before %s\n", e->role.c_str(), (t - t_prev) * 1e3, g->ops[(size_t)idx].name.c_str());
        t_prev = t;
      }
      const GOp &op = g->ops[(size_t)idx];
      if (op.plc_kind != "Host") fail(MB_ERR_UNIMPLEMENTED, op.name + ": only lowered graphs (@Host placements) can be replayed");
      const bool mine = !e->has_role || (!op.owners.empty() && op.owners[0] == e->role);
      Val &slot = e->values[(size_t)idx];
      if (op.kind == "Send") {
        if (!mine) continue;
        const Val &v = in(op, 0);
        const std::string *rk = op.attr("rendezvous_key"), *peer = op.attr("receiver");
        if (!rk || !peer) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": Send needs rendezvous_key and receiver");
        if (e->transport) send(v, unquote(*peer));
        else e->mailbox[*rk] = e->g->index.at(op.inputs[0]);
        release(c, slot);
        slot.kind = V_UNIT;
        e->executed++;
        if (e->transport) consumed(op);  // mailbox mode: the value stays until its Receive ran
        continue;
      }
      if (op.kind == "Receive") {
        if (!mine) continue;
        const std::string *rk = op.attr("rendezvous_key"), *peer = op.attr("sender");
        if (!rk || !peer) fail(MB_ERR_INVALID_ARGUMENT, op.name + ": Receive needs rendezvous_key and sender");
        Val v;
        if (e->transport) {
          v = recv(unquote(*peer));
        } else {
          auto it = e->mailbox.find(*rk);
          if (it == e->mailbox.end()) fail(MB_ERR_KERNEL, op.name + ": nothing was sent under this rendezvous key");
          const int src = it->second;
          v = copy(e->values[(size_t)src]);
          e->mailbox.erase(it);
          if (!e->keep_values && --uses[(size_t)src] <= 0) release(c, e->values[(size_t)src]);  // the Send's use
        }
        release(c, slot);
        slot = v;
        e->executed++;
        continue;
      }
      if (!mine) continue;
      Val v;
      auto ov = e->bound.find(op.name);
      if (ov != e->bound.end() && op.kind != "Input") v = copy(ov->second);
      else v = execute(op, idx);
      release(c, slot);
      slot = v;
      e->executed++;
      consumed(op);
    }
    if (e->transport) reap(true);
  }
};

int guard_parse(mb_ctx *ctx, const std::function<int()> &f) {
  try {
    return f();
  } catch (const ParseErr &p) {
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, p.msg);
  } catch (const ExecErr &x) {
    return mb_fail(ctx, x.code, x.msg);
  } catch (const std::exception &x) {
    return mb_fail(ctx, MB_ERR_KERNEL, x.what());
  }
}

}  // namespace

extern "C" int mb_graph_parse(mb_ctx *ctx, const char *text, mb_graph **out) {
  if (!text || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_graph_parse: NULL argument");
  return guard_parse(ctx, [&]() {
    std::unique_ptr<mb_graph> g(new mb_graph());
    const std::string all(text);
    size_t pos = 0;
    while (pos <= all.size()) {
      size_t nl = all.find('\n', pos);
      if (nl == std::string::npos) nl = all.size();
      const std::string line = trim(all.substr(pos, nl - pos));
      pos = nl + 1;
      if (line.empty() || line[0] == '#' || (line.size() > 1 && line[0] == '/' && line[1] == '/')) continue;
      GOp op = parse_line(line);
      if (g->index.count(op.name)) throw ParseErr{"operation " + op.name + " is defined twice"};
      g->index[op.name] = (int)g->ops.size();
      g->ops.push_back(std::move(op));
    }
    toposort(g.get());
    *out = g.release();
    return (int)MB_OK;
  });
}

extern "C" int mb_graph_free(mb_graph *g) {
  delete g;
  return MB_OK;
}

extern "C" int64_t mb_graph_num_ops(const mb_graph *g) { return g ? (int64_t)g->ops.size() : 0; }

// Name of the i-th operation in execution order (Receive right after its Send); NULL when out of range.
extern "C" const char *mb_graph_op_name(const mb_graph *g, int64_t i) {
  if (!g || i < 0 || i >= (int64_t)g->order.size()) return nullptr;
  return g->ops[(size_t)g->order[(size_t)i]].name.c_str();
}

extern "C" int mb_exec_create(mb_ctx *ctx, const mb_graph *g, const char *role, mb_exec **out) {
  if (!ctx || !g || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_exec_create: NULL argument");
  mb_exec *e = new mb_exec();
  e->ctx = ctx;
  e->g = g;
  e->has_role = role != nullptr;
  if (role) e->role = role;
  e->transport = nullptr;
  e->transport_user = nullptr;
  e->stream_ordered = false;
  e->values.resize(g->ops.size());
  e->executed = e->bytes_sent = 0;
  e->transport_seconds = e->sync_seconds = 0.0;
  e->keep_values = true;
  e->sends_reported = 0;
  *out = e;
  return MB_OK;
}

extern "C" int mb_exec_free(mb_exec *e) {
  if (!e) return MB_OK;
  for (auto &v : e->values) release(e->ctx, v);
  for (auto &kv : e->bound) release(e->ctx, kv.second);
  for (auto *t : e->in_flight) mb_tensor_free(e->ctx, t);
  delete e;
  return MB_OK;
}

extern "C" int mb_exec_set_transport(mb_exec *e, mb_transport_fn fn, void *user) {
  if (!e) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, "mb_exec_set_transport: NULL executor");
  e->transport = fn;
  e->transport_user = user;
  int64_t caps = 0;
  if (fn && fn(user, 5, "", &caps, sizeof(caps), 0) != 0) caps = 0;  // optional mode: an error means "none"
  e->stream_ordered = (caps & 1) != 0;
  return MB_OK;
}

extern "C" int mb_exec_set_keep_values(mb_exec *e, int keep) {
  if (!e) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, "mb_exec_set_keep_values: NULL executor");
  e->keep_values = keep != 0;
  return MB_OK;
}

extern "C" int mb_exec_bind_tensor(mb_exec *e, const char *name, const mb_tensor *t) {
  if (!e || !name || !t) return mb_fail(e ? e->ctx : nullptr, MB_ERR_INVALID_ARGUMENT, "mb_exec_bind_tensor: NULL argument");
  Val v;
  v.kind = V_TENSOR;
  v.t = mb_tensor_retain(const_cast<mb_tensor *>(t));
  auto it = e->bound.find(name);
  if (it != e->bound.end()) release(e->ctx, it->second);
  e->bound[name] = v;
  return MB_OK;
}

extern "C" int mb_exec_run(mb_exec *e) {
  if (!e) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, "mb_exec_run: NULL executor");
  return guard_parse(e->ctx, [&]() {
    MB_CUDA(e->ctx, cudaSetDevice(e->ctx->device));
    Runner(e).run();
    return (int)MB_OK;
  });
}

extern "C" uint64_t mb_exec_executed(const mb_exec *e) { return e ? e->executed : 0; }
extern "C" uint64_t mb_exec_bytes_sent(const mb_exec *e) { return e ? e->bytes_sent : 0; }
extern "C" double mb_exec_seconds(const mb_exec *e, int which) { return !e ? 0.0 : (which == 0 ? e->transport_seconds : e->sync_seconds); }

namespace {
int describe(mb_exec *e, int idx, int *kind, mb_tensor **tensor, uint8_t *bytes, size_t bytes_cap, size_t *nbytes, int64_t *shape, int *rank) {
  const Val &v = e->values[(size_t)idx];
  if (kind) *kind = v.kind;
  if (tensor) *tensor = v.kind == V_TENSOR ? mb_tensor_retain(v.t) : nullptr;
  const uint8_t *src = v.kind == V_BYTES ? v.bytes.data() : (const uint8_t *)v.str.data();
  const size_t n = v.kind == V_BYTES ? v.bytes.size() : (v.kind == V_STRING ? v.str.size() : 0);
  if (nbytes) *nbytes = n;
  if (bytes && n) memcpy(bytes, src, std::min(n, bytes_cap));
  if (rank) *rank = v.kind == V_SHAPE ? (int)v.shape.size() : 0;
  if (shape && v.kind == V_SHAPE)
    for (size_t k = 0; k < v.shape.size() && k < MB_MAX_RANK; k++) shape[k] = v.shape[k];
  return MB_OK;
}
}  // namespace

extern "C" int mb_exec_value(mb_exec *e, const char *op_name, int *kind, mb_tensor **tensor, uint8_t *bytes, size_t bytes_cap, size_t *nbytes,
                             int64_t *shape, int *rank) {
  if (!e || !op_name) return mb_fail(e ? e->ctx : nullptr, MB_ERR_INVALID_ARGUMENT, "mb_exec_value: NULL argument");
  auto it = e->g->index.find(op_name);
  if (it == e->g->index.end()) return mb_fail(e->ctx, MB_ERR_INVALID_ARGUMENT, std::string("mb_exec_value: no operation named ") + op_name);
  return describe(e, it->second, kind, tensor, bytes, bytes_cap, nbytes, shape, rank);
}

extern "C" int mb_exec_output(mb_exec *e, int saved, const char *key, mb_tensor **tensor) {
  if (!e || !key || !tensor) return mb_fail(e ? e->ctx : nullptr, MB_ERR_INVALID_ARGUMENT, "mb_exec_output: NULL argument");
  const auto &m = saved ? e->saved : e->outputs;
  auto it = m.find(key);
  if (it == m.end()) return mb_fail(e->ctx, MB_ERR_INVALID_ARGUMENT, std::string("mb_exec_output: nothing recorded under ") + key);
  const Val &v = e->values[(size_t)it->second];
  if (v.kind != V_TENSOR) return mb_fail(e->ctx, MB_ERR_TYPE_MISMATCH, std::string("mb_exec_output: ") + key + " is not a tensor");
  *tensor = mb_tensor_retain(v.t);
  return MB_OK;
}

// Keys recorded by Output (saved = 0) / Save (saved = 1) operations, in sorted order; NULL past the end.
extern "C" const char *mb_exec_output_key(const mb_exec *e, int saved, int64_t i) {
  if (!e || i < 0) return nullptr;
  const auto &m = saved ? e->saved : e->outputs;
  if (i >= (int64_t)m.size()) return nullptr;
  auto it = m.begin();
  std::advance(it, i);
  return it->first.c_str();
}
