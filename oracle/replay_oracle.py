"""TEST INFRASTRUCTURE: replays a parsed lowered computation on the oracle's host kernels, op for op, so every
intermediate value of the CUDA replay (moose_b200/executor.py) can be compared bit for bit."""
import numpy as np

from . import moose_oracle as mo

_RING_BITS = {"HostRing64Tensor": 64, "HostRing128Tensor": 128, "Ring64Tensor": 64, "Ring128Tensor": 128}


def replay(ops_in_order, sess: mo.Session, parse_constant, attr_hex16, arguments=None):
    vals, outputs, mailbox = {}, {}, {}
    arguments = arguments or {}
    for op in ops_in_order:
        x = [vals[i] for i in op.inputs]
        k = op.kind
        if k == "PrfKeyGen":
            v = sess.gen_key()
        elif k == "DeriveSeed":
            v = sess.derive_seed(attr_hex16(op, "sync_key"), x[0])
        elif k == "Constant":
            ty, val = parse_constant(op.attrs["value"])
            if ty in ("HostFloat64Tensor", "Float64Tensor"):
                v = np.array(val, dtype=np.float64)
            elif ty in _RING_BITS:
                v = mo.Ring.from_ints(_RING_BITS[ty], val)
            elif ty == "HostBitTensor":
                v = mo.Bits.from_array(np.array(val, dtype=np.uint8))
            else:
                v = val
        elif k == "Input":
            v = np.asarray(arguments[op.attrs["arg_name"].strip('"')], dtype=np.float64)
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
            v = mo.bit_decompose(x[0]) if op.output_type == "HostBitTensor" else mo.bit_decompose_ring(x[0])
        elif k == "RingInject":
            v = mo.ring_inject(_RING_BITS[op.output_type], int(op.attrs["bit_idx"]), x[0])
        elif k == "ShlDim":
            v = mo.shl_dim(int(op.attrs["amount"]), int(op.attrs["bit_length"]), x[0])
        elif k == "IndexAxis":
            v = mo.index_axis(int(op.attrs["axis"]), int(op.attrs["index"]), x[0])
        elif k == "Identity":
            v = x[0]
        elif k == "ExpandDims":
            v = mo.expand_dims([int(a) for a in op.attrs["axis"].strip("[] ").split(",") if a.strip()], x[0])
        elif k == "Squeeze":
            v = mo.squeeze(int(op.attrs["axis"]) if "axis" in op.attrs else None, x[0])
        elif k == "Transpose":
            v = mo.transpose(x[0])
        elif k == "Concat":
            v = mo.concat(int(op.attrs["axis"]), x)
        elif k == "Slice":
            f = {p.split("=")[0].strip(): int(p.split("=")[1]) for p in op.attrs["slice"].strip("{} ").split(",")}
            info = [(f["start"], f.get("end"), f.get("step"))]
            v = tuple(x[0][info[0][0]:info[0][1]]) if isinstance(x[0], tuple) else mo.slice_ring(info, x[0])
        elif k == "Diag":
            v = mo.diag(x[0])
        elif k == "Select":
            v = mo.select(int(op.attrs["axis"]), x[1], x[0])
        elif k == "Div":
            v = mo.ring_div(x[0], x[1])
        elif k == "Less":
            v = mo.host_less(x[0], x[1])
        elif k == "Greater":
            v = mo.host_greater(x[0], x[1])
        elif k == "Mux":
            v = mo.host_mux(x[0], x[1], x[2])
        elif k == "Sign":
            v = mo.host_sign(x[0])
        elif k == "Cast":
            v = mo.cast_ring128_to_ring64(x[0])
        elif k == "Maximum":
            v = mo.host_maximum(x)
        elif k == "Sum":
            axis = op.attrs.get("axis")
            v = mo.sum_axis(x[0], None if axis in (None, "None") else int(axis))
        elif k == "Or":
            v = mo.bit_op("or", x[0], x[1])
        elif k == "BitExtract":
            v = mo.bit_extract(int(op.attrs["bit_idx"]), x[0])
        elif k == "Reshape":
            v = x[0].reshape(x[1]) if isinstance(x[0], mo.Ring) else mo.Bits(x[1], x[0].packed)
        elif k == "Broadcast":
            v = mo.broadcast(x[1], x[0])
        elif k == "RingFixedpointMean":
            v = mo.fixedpoint_mean(int(op.attrs["axis"]) if "axis" in op.attrs else None, int(op.attrs["scaling_base"]),
                                   int(op.attrs["scaling_exp"]), x[0])
        elif k == "RingFixedpointArgmax":
            v = mo.fixedpoint_argmax(int(op.attrs["axis"]), x[0])
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
