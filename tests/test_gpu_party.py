"""One party per context (csrc/party.cpp): the three parties run as threads on one GPU and talk through the
in-process transport.  Every party has its own entropy, so results are checked by value (revealed outputs against the
plaintext computation, TruncPr within its 1-ulp error) and by structure (party i's second share is party i+1's
first share)."""
import threading

import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import BIT, RING64, RING128
from moose_b200.party_rt import LocalTransport, PartyMLP, PartySession, PeerMemory

pytestmark = pytest.mark.gpu
NAMES = ("player0", "player1", "player2")
DT = {64: RING64, 128: RING128}


_TRANSPORT = {"kind": "local"}


@pytest.fixture(autouse=True, params=["local", "peer"])
def transport_kind(request):
    """Every test runs over the in-process queues and over the peer-memory rings (csrc/peer_transport.cu; three
    contexts of one process on one GPU, 1 MiB rings so that payloads wrap and split into chunks)."""
    _TRANSPORT["kind"] = request.param
    yield request.param
    _TRANSPORT["kind"] = "local"


def run_parties(fn, session_id="party-test"):
    sessions = [mb.Session(0, session_id, bytes([17 * (i + 1)]) * 16) for i in range(3)]
    peers = []
    if _TRANSPORT["kind"] == "peer":
        peers = [PeerMemory(sessions[i], i, NAMES, 1 << 20) for i in range(3)]
        handles = [p.handle_bytes() for p in peers]
        for p in peers:
            p.connect(handles)
        endpoints = peers
    else:
        lt = LocalTransport(NAMES)
        endpoints = [lt.endpoint(NAMES[i], sessions[i]) for i in range(3)]
    results, errors = [None] * 3, []

    def worker(i):
        try:
            ps = PartySession(sessions[i], i, endpoints[i], NAMES)
            results[i] = fn(ps)
            ps.flush()
            ps.close()
        except Exception as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(3)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errors, errors
    assert all(not t.is_alive() for t in threads), "a party is stuck"
    for p in peers:
        p.close()
    for s in sessions:
        s.close()
    return results


def signed(a, bits):
    a = a.astype(object)
    if bits == 128:
        a = a[..., 0] + (a[..., 1] << 64)
    m = 1 << bits
    return np.where(a >= m // 2, a - m, a)


def ints(vals, bits, sess):
    flat = [int(v) % (1 << bits) for v in np.asarray(vals, dtype=object).reshape(-1)]
    shape = np.shape(vals)
    if bits == 64:
        return sess.from_host(RING64, np.array(flat, dtype=np.uint64).reshape(shape))
    arr = np.array([[v & (2 ** 64 - 1), v >> 64] for v in flat], dtype=np.uint64).reshape(shape + (2,))
    return sess.from_host(RING128, arr)


@pytest.mark.parametrize("bits", [64, 128])
def test_share_reveal_consistency_and_linear_ops(bits):
    rng = np.random.default_rng(bits)
    xv = rng.integers(-10 ** 6, 10 ** 6, (6, 5))
    yv = rng.integers(-10 ** 6, 10 ** 6, (6, 5))
    pub = rng.integers(-100, 100, (6, 5))

    def prog(p):
        s = p.sess
        x = p.share(0, ints(xv, bits, s) if p.me == 0 else None, DT[bits], xv.shape)
        y = p.share(2, ints(yv, bits, s) if p.me == 2 else None, DT[bits], yv.shape)
        z = p.add_public(p.sub(p.add(x, y), y), ints(pub, bits, s))
        m = p.mul(x, y)
        outs = [p.reveal(r, z) for r in range(3)]
        prod = p.reveal(1, m)
        return (x.share(0).to_host(), x.share(1).to_host(), m.share(0).to_host(), m.share(1).to_host(),
                [o.to_host() if o is not None else None for o in outs], prod.to_host() if prod is not None else None)

    res = run_parties(prog)
    for i in range(3):  # replicated structure: (x_i, x_{i+1})
        assert np.array_equal(res[i][1], res[(i + 1) % 3][0]) and np.array_equal(res[i][3], res[(i + 1) % 3][2])
        assert res[i][4][i] is not None and all(res[i][4][r] is None for r in range(3) if r != i)
        assert np.array_equal(signed(res[i][4][i], bits), xv + pub)
    assert np.array_equal(signed(res[1][5], bits), xv * yv)
    assert not np.array_equal(res[0][0], res[1][0])  # the shares are not the plaintext in disguise


@pytest.mark.parametrize("bits,shape", [(64, (130, 40, 24)), (128, (70, 33, 16)), (64, (256, 512, 256))])
def test_dot_and_trunc_pr(bits, shape):
    """Replicated dot (IMAD path and, for the large case, the two-term tensor-pipe GEMM per party) followed by
    TruncPr: fixed-point matrix product within the truncation error."""
    M, K, N = shape
    frac = 20
    rng = np.random.default_rng(M)
    a, b = rng.uniform(-2, 2, (M, K)), rng.uniform(-2, 2, (K, N))

    def prog(p):
        s = p.sess
        enc = lambda v: s.fixedpoint_encode(DT[bits], frac, s.from_host(mb.FLOAT64, v))
        x = p.share(0, enc(a) if p.me == 0 else None, DT[bits], a.shape)
        y = p.share(1, enc(b) if p.me == 1 else None, DT[bits], b.shape)
        z = p.trunc_pr(frac, p.dot(x, y))
        out = p.reveal(2, z)
        return s.fixedpoint_decode(frac, out).to_host() if out is not None else None

    res = run_parties(prog)
    assert res[0] is None and res[1] is None
    assert np.abs(res[2] - a @ b).max() < K * 2.0 ** -frac * 4 + 2.0 ** -(frac - 2)


@pytest.mark.parametrize("bits", [64, 128])
def test_msb_ring_inject_and_maximum(bits):
    rng = np.random.default_rng(5 + bits)
    xv = rng.integers(-2 ** 40, 2 ** 40, (8, 32))
    xv[0, :4] = [0, -1, 1, -(2 ** 62)]
    yv = rng.integers(-2 ** 40, 2 ** 40, (8, 32))

    def prog(p):
        s = p.sess
        x = p.share(1, ints(xv, bits, s) if p.me == 1 else None, DT[bits], xv.shape)
        y = p.share(0, ints(yv, bits, s) if p.me == 0 else None, DT[bits], yv.shape)
        m = p.msb(x)
        inj = p.ring_inject(DT[bits], 3, m)
        mx = p.maximum([x, y])
        lt = p.less(x, y)
        r = [p.reveal(0, m), p.reveal(0, inj), p.reveal(0, mx), p.reveal(0, lt)]
        return [t.bits() if t is not None and t.dtype == BIT else (t.to_host() if t is not None else None) for t in r]

    res = run_parties(prog)[0]
    assert np.array_equal(res[0], (xv < 0).astype(np.uint8))
    assert np.array_equal(signed(res[1], bits), (xv < 0).astype(object) * 8)
    assert np.array_equal(signed(res[2], bits), np.maximum(xv, yv))
    assert np.array_equal(res[3], (xv < yv).astype(np.uint8))


@pytest.mark.parametrize("bits", [64, 128])
def test_ragged_sizes_mul_trunc_inject(bits):
    """Element counts that fill neither a 16-byte vector (odd, Ring64) nor a 32-bit word of a bit tensor: the fused
    party steps mask the padding (opened top bit of TruncPr packed by ballot, dabit bits read per element)."""
    rng = np.random.default_rng(bits + 1)
    shape = (3, 5)
    xv, yv = rng.integers(-2 ** 30, 2 ** 30, shape), rng.integers(-2 ** 20, 2 ** 20, shape)
    bv = rng.integers(0, 2, shape).astype(np.uint8)

    def prog(p):
        s = p.sess
        x = p.share(0, ints(xv, bits, s) if p.me == 0 else None, DT[bits], shape)
        y = p.share(2, ints(yv, bits, s) if p.me == 2 else None, DT[bits], shape)
        b = p.share(1, s.from_bits(bv) if p.me == 1 else None, BIT, shape)
        m = p.mul(x, y)
        t = p.trunc_pr(7, m)
        inj = p.ring_inject(DT[bits], 2, b)
        r = [p.reveal(1, m), p.reveal(1, t), p.reveal(1, inj)]
        return [v.to_host() if v is not None else None for v in r]

    res = run_parties(prog)[1]
    prod = xv.astype(object) * yv.astype(object)
    assert np.array_equal(signed(res[0], bits), prod)
    assert np.abs(signed(res[1], bits) - np.array([[int(v) >> 7 for v in row] for row in prod], dtype=object)).max() <= 1
    assert np.array_equal(signed(res[2], bits), bv.astype(object) * 4)


def test_party_mlp_matches_float():
    rng = np.random.default_rng(9)
    dims = (24, 16, 16, 4)
    ws = [rng.normal(0, dims[i] ** -0.5, (dims[i], dims[i + 1])) for i in range(3)]
    bs = [rng.normal(0, 0.1, dims[i + 1]) for i in range(3)]
    x = rng.uniform(0, 1, (32, dims[0]))

    def prog(p):
        mlp = PartyMLP(p, ws, bs, RING64, 23)
        s = p.sess
        xf = s.fixedpoint_encode(RING64, 23, s.from_host(mb.FLOAT64, x)) if p.me == 0 else None
        out = mlp.forward(xf, x.shape[0], reveal_to=0)
        return out.to_host() if out is not None else None, p.bytes_sent

    res = run_parties(prog)
    h = x
    for i in range(3):
        h = h @ ws[i] + bs[i]
        if i < 2:
            h = np.maximum(h, 0)
    assert np.abs(res[0][0] - h).max() < 1e-4 and res[1][0] is None
    assert all(r[1] > 0 for r in res)


def test_lowered_graph_three_executors_over_peer_memory(transport_kind):
    """The native executor in party-per-process mode over the peer-memory transport: the lowered + networked MLP
    graph (moose_b200/lowering.py), one executor per party as threads on one GPU, every Send / Receive through the
    rings (stream-ordered: no host synchronisation per transfer)."""
    if transport_kind != "peer":
        pytest.skip("one transport is enough here")
    from moose_b200.executor import Executor
    from moose_b200.lowering import Lowering, networking_pass, prune_pass

    dims, batch = (6, 5, 4, 3), 4
    rng = np.random.default_rng(0)
    x = rng.uniform(0, 1, (batch, dims[0]))
    ws = [rng.normal(0, dims[l] ** -0.5, (dims[l], dims[l + 1])) for l in range(3)]
    bs = [rng.normal(0, 0.1, (1, dims[l + 1])) for l in range(3)]
    want = x
    for l in range(3):
        want = want @ ws[l] + bs[l]
        if l < 2:
            want = np.maximum(want, 0)
    net = networking_pass(prune_pass(Lowering(64).mlp_program(3, 23)))
    sessions = [mb.Session(0, "peer-exec", None) for _ in range(3)]
    peers = [PeerMemory(sessions[i], i, NAMES, 1 << 20) for i in range(3)]
    handles = [p.handle_bytes() for p in peers]
    for p in peers:
        p.connect(handles)
    results, errors = [None] * 3, []

    def worker(i):
        try:
            s = sessions[i]
            over = {}
            for l in range(3):
                over[f"w{l}"], over[f"b{l}"] = s.from_host(mb.FLOAT64, ws[l]), s.from_host(mb.FLOAT64, bs[l])
            if i == 0:
                over["x"] = s.from_host(mb.FLOAT64, x)
            ex = Executor(s, role=NAMES[i], comm=peers[i], overrides=over, keep_values=False)
            outs = ex.run(net)
            results[i] = {k: v.to_host() for k, v in outs.items()}
            s.sync()
            ex.close()
        except Exception as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(3)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=200)
    assert not errors, errors
    assert all(not t.is_alive() for t in threads), "a party is stuck"
    for p in peers:
        p.close()
    for s in sessions:
        s.close()
    assert np.abs(results[0]["output_0"] - want).max() < 1e-5
