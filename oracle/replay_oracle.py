"""TEST INFRASTRUCTURE: replays a parsed lowered computation on the oracle's host kernels, op for op, so every
intermediate value of the CUDA replay (moose_b200/executor.py) can be compared bit for bit."""
import numpy as np

from . import moose_oracle as mo

_RING_BITS = {"HostRing64Tensor": 64, "HostRing128Tensor": 128, "Ring64Tensor": 64, "Ring128Tensor": 128}


def replay(ops_in_order, sess: mo.Session, parse_constant, attr_hex16):
    vals, outputs, mailbox = {}, {}, {}
    for op in ops_in_order:
        x = [vals[i] for i in op.inputs]
        k = op.kind
        if k == "PrfKeyGen":
            v = sess.gen_key()
        elif k == "DeriveSeed":
            v = sess.derive_seed(attr_hex16(op, "sync_key"), x[0])
        elif k == "Constant":
            ty, val = parse_constant(op.attrs["value"])
            v = np.array(val, dtype=np.float64) if ty in ("HostFloat64Tensor", "Float64Tensor") else val
        elif k == "RingFixedpointEncode":
            v = mo.fixed_encode(_RING_BITS[op.output_type], x[0], int(op.attrs["scaling_exp"]))
        elif k == "RingFixedpointDecode":
            v = mo.fixed_decode(x[0], int(op.attrs["scaling_exp"]))
        elif k == "Shape":
            v = tuple(x[0].shape)
        elif k == "Fill":
            _, val = parse_constant(op.attrs["value"])
            v = mo.fill_bits(int(val), x[0]) if op.output_type == "HostBitTensor" else mo.fill(_RING_BITS[op.output_type], int(val), x[0])
        elif k in ("SampleSeeded", "Sample"):
            seed = x[1] if k == "SampleSeeded" else sess.entropy.next16()
            if op.output_type == "HostBitTensor":
                v = mo.sample_uniform_seeded_bits(x[0], seed, sess.bit_policy)
            elif "max_value" in op.attrs:
                v = mo.sample_bits_seeded_ring(_RING_BITS[op.output_type], x[0], seed, sess.bit_policy)
            else:
                v = mo.sample_uniform_seeded(_RING_BITS[op.output_type], x[0], seed)
        elif k == "Add":
            v = mo.add(x[0], x[1])
        elif k == "Sub":
            v = mo.sub(x[0], x[1])
        elif k == "Mul":
            v = mo.mul(x[0], x[1])
        elif k == "Dot":
            v = mo.dot(x[0], x[1])
        elif k == "Neg":
            v = mo.neg(x[0])
        elif k == "Shl":
            v = mo.shl(int(op.attrs["amount"]), x[0])
        elif k == "Shr":
            v = mo.shr(int(op.attrs["amount"]), x[0])
        elif k == "AddN":
            v = mo.add_n(x)
        elif k == "Xor":
            v = mo.bit_op("xor", x[0], x[1])
        elif k == "And":
            v = mo.bit_op("and", x[0], x[1]) if isinstance(x[0], mo.Bits) else mo.ring_and(x[0], x[1])
        elif k == "BitDecompose":
            v = mo.bit_decompose(x[0])
        elif k == "RingInject":
            v = mo.ring_inject(_RING_BITS[op.output_type], int(op.attrs["bit_idx"]), x[0])
        elif k == "ShlDim":
            v = mo.shl_dim(int(op.attrs["amount"]), int(op.attrs["bit_length"]), x[0])
        elif k == "IndexAxis":
            assert int(op.attrs["axis"]) == 0
            v = mo.index_axis0(int(op.attrs["index"]), x[0])
        elif k == "Send":
            mailbox[op.attrs["rendezvous_key"]] = x[0]
            v = None
        elif k == "Receive":
            v = mailbox.pop(op.attrs["rendezvous_key"])
        elif k == "Output":
            outputs[op.attrs["tag"].strip('"')] = x[0]
            v = x[0]
        elif k == "Save":
            outputs["saved:" + str(x[0])] = x[1]
            v = None
        else:
            raise NotImplementedError(k)
        vals[op.name] = v
    return vals, outputs
