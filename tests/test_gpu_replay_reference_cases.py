"""The reference's executor tests on ring graphs (moose/src/execution/mod.rs:200-222 seeded sample, :741-771 ring
binops, :773-846 ring dots, :848-918 fill, :920-956 shr), written as the reference writes them -- nullary operations
without an argument list -- and run through the native executor on the GPU; every value against the oracle replay."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import textual as tx
from moose_b200.executor import Executor, toposort
from oracle import moose_oracle as mo
from oracle import replay_oracle as ro
from _util import ENTROPY_KEY
from test_gpu_replay import same_value

pytestmark = pytest.mark.gpu


def run_both(source):
    text = "\n".join(line.strip() for line in source.strip().splitlines()) + "\n"
    ops = tx.parse_computation(text)
    sess = mb.Session(0, "foobar", ENTROPY_KEY)
    try:
        ex = Executor(sess)
        ex.run(text)                      # the native reader parses the text itself
        assert ex.executed == len(ops)
        ovals, _ = ro.replay(toposort(ops), mo.Session("foobar", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16)
        for op in ops:
            assert same_value(ex.values[op.name], ovals[op.name]), op.name
        out = {}
        for op in ops:                    # fetched before the session closes
            v = ex.values[op.name]
            out[op.name] = (v.shape, v.to_host()) if isinstance(v, mb.Tensor) else v
        return out
    finally:
        sess.close()


@pytest.mark.parametrize("op,expected", [("Add", [5]), ("Mul", [6]), ("Sub", [1])])
def test_ring_binop_graphs(op, expected):
    vals = run_both(f"""
        x = Constant{{value=HostRing64Tensor([3])}}: () -> HostRing64Tensor @Host(alice)
        y = Constant{{value=HostRing64Tensor([2])}}: () -> HostRing64Tensor @Host(alice)
        res = {op} : (HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor (x, y) @Host(alice)
        output = Output{{tag = "output_0"}} : (HostRing64Tensor) -> HostRing64Tensor (res) @Host(alice)""")
    assert vals["res"][1].tolist() == expected


@pytest.mark.parametrize("x,y,expected", [
    ("[[1, 2], [3, 4]]", "[[1, 0], [0, 1]]", [[1, 2], [3, 4]]),
    ("[[1, 2], [3, 4]]", "[1, 1]", [3, 7]),
    ("[1, 1]", "[[1, 2], [3, 4]]", [4, 6]),
])
def test_ring_dot_graphs(x, y, expected):
    vals = run_both(f"""
        x = Constant{{value=HostRing64Tensor({x})}}: () -> HostRing64Tensor @Host(alice)
        y = Constant{{value=HostRing64Tensor({y})}}: () -> HostRing64Tensor @Host(alice)
        res = Dot : (HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor (x, y) @Host(alice)
        output = Output{{tag = "output_0"}} : (HostRing64Tensor) -> HostRing64Tensor (res) @Host(alice)""")
    assert vals["res"][1].tolist() == expected


@pytest.mark.parametrize("ty,shape", [("Ring64", "2"), ("Ring128", "2"), ("Ring64", "2, 1"), ("Ring128", "2, 3")])
def test_fill_shr_and_sample_graphs(ty, shape):
    dims = tuple(int(d) for d in shape.split(","))
    vals = run_both(f"""
        shape = Constant{{value=HostShape([{shape}])}}: () -> HostShape @Host(alice)
        res = Fill {{value = {ty}(1)}} : (HostShape) -> Host{ty}Tensor (shape) @Host(alice)
        four = Fill {{value = {ty}(4)}} : (HostShape) -> Host{ty}Tensor (shape) @Host(alice)
        half = Shr {{amount = 1}}: (Host{ty}Tensor) -> Host{ty}Tensor (four) @Host(alice)
        seed = Constant{{value=HostSeed(00000000000000000000000000000000)}}: () -> HostSeed @Host(alice)
        sampled = SampleSeeded{{}}: (HostShape, HostSeed) -> Host{ty}Tensor (shape, seed) @Host(alice)
        sshape = Shape: (Host{ty}Tensor) -> HostShape (sampled) @Host(alice)
        output = Output{{tag = "output_0"}} : (Host{ty}Tensor) -> Host{ty}Tensor (res) @Host(alice)""")
    assert vals["res"][0] == dims and tuple(vals["sshape"]) == dims
    words = 1 if ty == "Ring64" else 2
    half = vals["half"][1].reshape(-1, words)
    assert half[:, 0].tolist() == [2] * int(np.prod(dims))
