"""CPU tests of the host-dialect textual reader, the replay order and the oracle-side replay
(moose/src/textual/parsing.rs grammar; lowered exemplar tutorials/dotprod-compiled.moose)."""
import glob
import gzip
import os

import numpy as np
import pytest

from moose_b200 import textual as tx
from moose_b200.executor import toposort
from oracle import moose_oracle as mo
from oracle import replay_oracle as ro

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return tx.loads_json(gzip.open(os.path.join(GOLDEN, name + ".json.gz")).read().decode())


def test_parse_host_dialect_lines():
    text = '''
    k = PrfKeyGen: () -> HostPrfKey () @Host(carole)
    s = DeriveSeed{sync_key = 0102030405060708090a0b0c0d0e0f10}: (HostPrfKey) -> HostSeed (k) @Host(carole)
    c = Constant{value = HostFloat64Tensor([[1.5, -2.0], [0.25, 4.0]])}: () -> HostFloat64Tensor () @Host(alice)
    e = RingFixedpointEncode{scaling_base = 2, scaling_exp = 23}: (HostFloat64Tensor) -> HostRing64Tensor (c) @Host(alice)
    n = AddN: [HostRing64Tensor] -> HostRing64Tensor (e, e, e) @Host(alice)
    snd = Send{rendezvous_key = 07000000000000000000000000000000, receiver = "bob"}: (HostRing64Tensor) -> HostUnit (n) @Host(alice)
    rcv = Receive{rendezvous_key = 07000000000000000000000000000000, sender = "alice"}: () -> HostRing64Tensor () @Host(bob)
    o = Output{tag = "out, with (brackets)"}: (HostRing64Tensor) -> HostRing64Tensor (rcv) @Host(bob)
    '''
    ops = tx.parse_computation(text)
    assert [o.kind for o in ops] == ["PrfKeyGen", "DeriveSeed", "Constant", "RingFixedpointEncode", "AddN", "Send",
                                     "Receive", "Output"]
    assert ops[1].inputs == ["k"] and tx.attr_hex16(ops[1], "sync_key") == bytes(range(1, 17))
    assert tx.parse_constant(ops[2].attrs["value"]) == ("HostFloat64Tensor", [[1.5, -2.0], [0.25, 4.0]])
    assert ops[4].variadic and ops[4].inputs == ["e", "e", "e"] and ops[4].input_types == ["HostRing64Tensor"]
    assert tx.attr_str(ops[5], "receiver") == "bob" and ops[6].owners == ["bob"]
    assert tx.attr_str(ops[7], "tag") == "out, with (brackets)"
    assert tx.loads_json(tx.dumps_json(ops)) == ops


def test_parse_logical_dialect_types_and_placements():
    op = tx.parse_line('dot_0 = Dot: (Tensor<Fixed128(24, 40)>, Tensor<Fixed128(24, 40)>) -> Tensor<Fixed128(24, 40)> '
                       '(cast_0, cast_1) @Replicated(player0, player1, player2)')
    assert op.input_types == ["Tensor<Fixed128(24, 40)>"] * 2 and op.output_type == "Tensor<Fixed128(24, 40)>"
    assert op.placement_kind == "Replicated" and op.owners == ["player0", "player1", "player2"]
    with pytest.raises(tx.ParseError):
        tx.parse_line("x = Add: (A, B) -> C (a, b)")  # no placement


def test_sync_key_byte_list_form():
    """execution/mod.rs:176: `DeriveSeed {sync_key = [1, 2, 3]}`."""
    op = tx.parse_line("seed = DeriveSeed {sync_key = [1, 2, 3]}: (HostPrfKey) -> HostSeed (key) @Host(alice)")
    assert tx.attr_hex16(op, "sync_key") == bytes([1, 2, 3]) + bytes(13)


def test_toposort_keeps_receive_after_send():
    ops = load("dotprod_networked")
    order = toposort(ops)
    idx = {op.name: i for i, op in enumerate(order)}
    for op in order:
        for i in op.inputs:
            assert idx[i] < idx[op.name]
    sends = {op.attrs["rendezvous_key"]: idx[op.name] for op in order if op.kind == "Send"}
    for op in order:
        if op.kind == "Receive":
            assert idx[op.name] == sends[op.attrs["rendezvous_key"]] + 1


@pytest.mark.parametrize("name,tag,expected", [("dotprod_compiled", "output_0", 32.0), ("dotprod_networked", "output_0", 32.0),
                                               ("rep_computation", "saved:y_uri", np.exp(2.0))])
def test_oracle_replays_reference_lowered_graphs(name, tag, expected):
    """The reference's own compiled graphs -- tutorials/dotprod-compiled.moose (x.w = 32.0, BASELINE configs[0]) and
    the 19 045-op moose/benches/rep_computation.moose (exp(2.0) on Fixed128(8,27)) -- run on the oracle's host
    kernels: an end-to-end pin of the kernels' semantics against the reference's op sequences."""
    order = toposort(load(name))
    _, outs = ro.replay(order, mo.Session("replay", bytes(range(16))), tx.parse_constant, tx.attr_hex16)
    assert abs(float(np.asarray(outs[tag]).reshape(-1)[0]) - expected) < 1e-5


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree only exists in the build container")
def test_reader_on_every_reference_moose_file_and_fixtures_are_current():
    files = glob.glob("/root/reference/**/*.moose", recursive=True)
    assert len(files) >= 6
    for f in files:
        assert len(tx.parse_computation(open(f).read())) > 0
    assert load("dotprod_compiled") == tx.parse_computation(open("/root/reference/tutorials/dotprod-compiled.moose").read())


def test_host_tail_graph_on_oracle():
    """Shape / indexing / plaintext-helper operators: parse a host-dialect graph written for these tests and
    replay it on the oracle; plain values checked with numpy."""
    from _util import ENTROPY_KEY, HOST_TAIL_GRAPH
    ops = tx.parse_computation(HOST_TAIL_GRAPH)
    assert len(ops) == 33 and tx.attr_slice(ops[11], "slice") == [(0, None, -1)] and tx.attr_int_list(ops[4], "axis") == [1]
    vals, outs = ro.replay(toposort(ops), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16)
    x = np.array([[1.5, -2.0, 3.25], [0.5, 4.0, -1.0]])
    cat = np.concatenate([x[:, :1], x], axis=1)
    assert np.array_equal(outs["mean"], cat.mean(axis=0)) and np.array_equal(outs["abs"], np.abs(x))
    assert vals["amax"].to_ints().tolist() == cat.argmax(axis=1).tolist()
    assert vals["tail"] == (4,) and vals["sel"].shape == (3, 2) and vals["planes"].shape == (64, 2)
    assert vals["dg"].to_ints().tolist() == [int(1.5 * 65536), int(4.0 * 65536)]
    assert vals["either"].to_array().tolist() == [[1, 1, 1], [1, 1, 1]] and vals["nlt"].to_array().tolist() == [[1, 0, 1], [1, 1, 0]]
    assert vals["total"].to_ints().tolist() == int(np.abs(x).sum() * 65536)


def test_native_parser_matches_python_order():
    """The C++ reader / scheduler of libmoose_b200 (csrc/executor.cpp) against the Python one: same operation count
    and the same execution order (Receive directly after its Send) on every fixture.  Parsing needs no GPU."""
    import ctypes
    import moose_b200 as mb
    from _util import HOST_TAIL_GRAPH
    lib = mb.lib()
    graphs = [load(n) for n in ("dotprod_compiled", "dotprod_networked", "rep_computation")] + [tx.parse_computation(HOST_TAIL_GRAPH)]
    for ops in graphs:
        g = ctypes.c_void_p()
        assert lib.mb_graph_parse(None, tx.dumps_text(ops).encode(), ctypes.byref(g)) == 0
        n = lib.mb_graph_num_ops(g)
        assert n == len(ops)
        native = [lib.mb_graph_op_name(g, i).decode() for i in range(n)]
        assert native == [op.name for op in toposort(ops)]
        lib.mb_graph_free(g)
    g = ctypes.c_void_p()
    assert lib.mb_graph_parse(None, b"x = Add (a, b) @Host(alice)\n", ctypes.byref(g)) == -1
    assert lib.mb_graph_parse(None, b"x = Add: (T, T) -> T (a, x) @Host(alice)\n", ctypes.byref(g)) == -1  # cycle
    # a nullary operation may leave out its "()" (textual/parsing.rs:176), as the reference's executor tests write them
    src = ("x = Constant{value=HostRing64Tensor([3])}: () -> HostRing64Tensor @Host(alice)\n"
           "s = Shape: (HostRing64Tensor) -> HostShape (x) @Host(alice)\n"
           "t = Mul: (Tensor<Fixed128(24, 40)>, Tensor<Fixed128(24, 40)>) -> Tensor<Fixed128(24, 40)> (x, x) @Replicated(alice, bob, carole)\n")
    assert lib.mb_graph_parse(None, src.encode(), ctypes.byref(g)) == 0 and lib.mb_graph_num_ops(g) == 3
    lib.mb_graph_free(g)
    ops = tx.parse_computation(src)
    assert ops[0].inputs == [] and ops[1].inputs == ["x"] and ops[2].output_type == "Tensor<Fixed128(24, 40)>"
    assert ops[2].inputs == ["x", "x"] and ops[2].owners == ["alice", "bob", "carole"]
    with pytest.raises(tx.ParseError):
        tx.parse_line("x = Add: (T, T) -> T junk @Host(alice)")


def _replay_text(source):
    ops = tx.parse_computation("\n".join(line.strip() for line in source.strip().splitlines()))
    return ro.replay(toposort(ops), mo.Session("foobar", bytes(range(16))), tx.parse_constant, tx.attr_hex16)


@pytest.mark.parametrize("op,expected", [("Add", [5]), ("Mul", [6]), ("Sub", [1])])
def test_reference_ring_binop_graphs(op, expected):
    """moose/src/execution/mod.rs:741-771 (test_ring_binop_invocation)."""
    vals, _ = _replay_text(f"""
        x = Constant{{value=HostRing64Tensor([3])}}: () -> HostRing64Tensor @Host(alice)
        y = Constant{{value=HostRing64Tensor([2])}}: () -> HostRing64Tensor @Host(alice)
        res = {op} : (HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor (x, y) @Host(alice)
        output = Output{{tag = "output_0"}} : (HostRing64Tensor) -> HostRing64Tensor (res) @Host(alice)""")
    assert vals["res"].to_ints().tolist() == expected


@pytest.mark.parametrize("x,y,expected", [
    ("[[1, 2], [3, 4]]", "[[1, 0], [0, 1]]", [[1, 2], [3, 4]]),
    ("[[1, 2], [3, 4]]", "[1, 1]", [3, 7]),
    ("[1, 1]", "[[1, 2], [3, 4]]", [4, 6]),
])
def test_reference_ring_dot_graphs(x, y, expected):
    """moose/src/execution/mod.rs:773-846 (test_ring_dot_invocation)."""
    vals, _ = _replay_text(f"""
        x = Constant{{value=HostRing64Tensor({x})}}: () -> HostRing64Tensor @Host(alice)
        y = Constant{{value=HostRing64Tensor({y})}}: () -> HostRing64Tensor @Host(alice)
        res = Dot : (HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor (x, y) @Host(alice)
        output = Output{{tag = "output_0"}} : (HostRing64Tensor) -> HostRing64Tensor (res) @Host(alice)""")
    assert vals["res"].to_ints().tolist() == expected


@pytest.mark.parametrize("ty,shape", [("Ring64", "2"), ("Ring128", "2"), ("Ring64", "2, 1"), ("Ring64", "2, 2"),
                                      ("Ring64", "1, 2"), ("Ring128", "2, 3")])
def test_reference_fill_graphs(ty, shape):
    """moose/src/execution/mod.rs:848-918 (test_fill)."""
    vals, _ = _replay_text(f"""
        shape = Constant{{value=HostShape([{shape}])}}: () -> HostShape @Host(alice)
        res = Fill {{value = {ty}(1)}} : (HostShape) -> Host{ty}Tensor (shape) @Host(alice)
        output = Output{{tag = "output_0"}} : (Host{ty}Tensor) -> Host{ty}Tensor (res) @Host(alice)""")
    dims = tuple(int(d) for d in shape.split(","))
    assert vals["res"].bits == int(ty[4:]) and vals["res"].shape == dims
    assert vals["res"].to_ints().reshape(-1).tolist() == [1] * int(np.prod(dims))


@pytest.mark.parametrize("ty", ["HostRing64Tensor", "HostRing128Tensor"])
def test_reference_shr_and_sample_graphs(ty):
    """moose/src/execution/mod.rs:920-956 (test_ring_bitwise_ops) and :200-222 (test_constants_sample_ring)."""
    vals, _ = _replay_text("""
        x = Constant{value=HostRing64Tensor([4, 4])}: () -> HostRing64Tensor @Host(alice)
        res = Shr {amount = 1}: (HostRing64Tensor) -> HostRing64Tensor (x) @Host(alice)
        output = Output{tag = "output_0"}: (HostRing64Tensor) -> HostRing64Tensor (res) @Host(alice)""".replace("HostRing64Tensor", ty))
    assert vals["res"].to_ints().tolist() == [2, 2]
    vals, _ = _replay_text("""
        seed = Constant{value=HostSeed(00000000000000000000000000000000)}: () -> HostSeed @Host(alice)
        xshape = Constant{value=HostShape([2, 2])}: () -> HostShape @Host(alice)
        sampled = SampleSeeded{}: (HostShape, HostSeed) -> HostRing64Tensor (xshape, seed) @Host(alice)
        shape = Shape: (HostRing64Tensor) -> HostShape (sampled) @Host(alice)
        output = Output{tag = "output_0"}: (HostShape) -> HostShape (shape) @Host(alice)""".replace("HostRing64Tensor", ty))
    assert tuple(vals["shape"]) == (2, 2)
    want = mo.sample_uniform_seeded(int(ty[8:-6]), (2, 2), bytes(16))
    assert vals["sampled"] == want


def test_random_graphs_parse_and_schedule_alike_in_python_and_native():
    """Differential test of the two readers / schedulers on seeded random DAGs: operations in shuffled textual order,
    attributes with nested brackets and quoted strings, nullary operations with and without "()", variadic
    signatures, parametrised types, Send / Receive pairs.  Same operation count, same execution order, and the text
    round-trips through the Python reader."""
    import ctypes
    import random
    import moose_b200 as mb
    lib = mb.lib()
    roles = ["alice", "bob", "carole"]
    for seed in range(25):
        rnd = random.Random(seed)
        lines, names = [], []
        for i in range(rnd.randint(3, 40)):
            name, role = f"op_{seed}_{i}", rnd.choice(roles)
            k = rnd.randint(0, min(3, len(names)))
            ins = rnd.sample(names, k)
            ty = rnd.choice(["HostRing64Tensor", "HostRing128Tensor", "HostBitTensor", "Tensor<Fixed128(24, 40)>"])
            if k == 0:
                attrs = rnd.choice(['{value = HostShape([2, 3])}', '{value=HostString("a, (b) = ]c")}', ""])
                arglist = rnd.choice([" ()", ""])
                lines.append(f"{name} = Constant{attrs}: () -> {ty}{arglist} @Host({role})")
            elif rnd.random() < 0.25:
                sig = f"[{ty}]"
                lines.append(f"{name} = AddN: {sig} -> {ty} ({', '.join(ins)}) @Host({role})")
            elif rnd.random() < 0.3:
                key = "%032x" % rnd.getrandbits(128)
                peer = rnd.choice([r for r in roles if r != role])
                lines.append(f'{name} = Send{{rendezvous_key = {key}, receiver = "{peer}"}}: ({ty}) -> HostUnit ({ins[0]}) @Host({role})')
                names.append(name)
                name = name + "_r"
                lines.append(f'{name} = Receive{{rendezvous_key = {key}, sender = "{role}"}}: () -> {ty} () @Host({peer})')
            else:
                attrs = rnd.choice(["", "{amount = 3}", "{slice = {start = 1, end = 4, step = 2}}", "{axis = [0, 1]}"])
                sig = "(" + ", ".join([ty] * k) + ")"
                lines.append(f"{name} = Op{i % 7}{attrs}: {sig} -> {ty} ({', '.join(ins)}) @Host({role})")
            names.append(name)
        rnd.shuffle(lines)
        text = "\n".join(lines) + "\n"
        ops = tx.parse_computation(text)
        assert len(ops) == len(lines)
        assert tx.parse_computation(tx.dumps_text(ops)) == ops
        g = ctypes.c_void_p()
        assert lib.mb_graph_parse(None, text.encode(), ctypes.byref(g)) == 0, (seed, lib.mb_last_error(None))
        n = lib.mb_graph_num_ops(g)
        native = [lib.mb_graph_op_name(g, i).decode() for i in range(n)]
        lib.mb_graph_free(g)
        assert native == [op.name for op in toposort(ops)], seed
