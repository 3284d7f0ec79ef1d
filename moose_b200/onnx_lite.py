"""Minimal reader for the parts of an ONNX file the tree-ensemble predictors need (no onnx package in this image):
walks the protobuf wire format of ModelProto -> graph -> node[op_type] -> attribute and graph.input[0] shape.
Field numbers from the public onnx.proto: ModelProto.graph = 7; GraphProto.node = 1, input = 11; NodeProto.op_type = 4,
attribute = 5; AttributeProto.name = 1, f = 2, i = 3, s = 4, floats = 7, ints = 8, strings = 9, type = 20;
ValueInfoProto.type = 2; TypeProto.tensor_type = 1; Tensor.shape = 2; TensorShapeProto.dim = 1; Dimension.dim_value = 1."""
import struct


def _varint(b, i):
    v, s = 0, 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << s
        s += 7
        if not c & 0x80:
            return v, i


def fields(b):
    """[(field number, wire type, value)] of one message; length-delimited values stay bytes."""
    out, i = [], 0
    while i < len(b):
        key, i = _varint(b, i)
        fn, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(b, i)
        elif wt == 1:
            v, i = b[i:i + 8], i + 8
        elif wt == 2:
            n, i = _varint(b, i)
            v, i = b[i:i + n], i + n
        elif wt == 5:
            v, i = b[i:i + 4], i + 4
        else:
            raise ValueError(f"wire type {wt}")
        out.append((fn, wt, v))
    return out


def _signed(v):
    return v - (1 << 64) if v >= 1 << 63 else v


def _attribute(b):
    name, a = None, {"floats": [], "ints": [], "strings": []}
    for fn, wt, v in fields(b):
        if fn == 1:
            name = v.decode()
        elif fn == 2:
            a["f"] = struct.unpack("<f", v)[0]
        elif fn == 3:
            a["i"] = _signed(v)
        elif fn == 4:
            a["s"] = v.decode()
        elif fn == 7:
            a["floats"] += list(struct.unpack(f"<{len(v) // 4}f", v)) if wt == 2 else [struct.unpack("<f", v)[0]]
        elif fn == 8:
            if wt == 2:
                i = 0
                while i < len(v):
                    x, i = _varint(v, i)
                    a["ints"].append(_signed(x))
            else:
                a["ints"].append(_signed(v))
        elif fn == 9:
            a["strings"].append(v.decode())
    return name, a


def find_node(model_bytes, op_type):
    """Attributes {name: {...}} of the first node with this op_type, and the feature count of graph.input[0]."""
    graph = next(v for fn, _, v in fields(model_bytes) if fn == 7)
    n_features, node_attrs = None, None
    for fn, _, v in fields(graph):
        if fn == 1 and node_attrs is None:
            fs = fields(v)
            if any(f == 4 and val.decode() == op_type for f, _, val in fs):
                node_attrs = dict(_attribute(val) for f, _, val in fs if f == 5)
        elif fn == 11 and n_features is None:
            ty = next(val for f, _, val in fields(v) if f == 2)
            tensor = next(val for f, _, val in fields(ty) if f == 1)
            shape = next(val for f, _, val in fields(tensor) if f == 2)
            dims = []
            for f, _, d in fields(shape):
                if f == 1:
                    dv = [val for g, _, val in fields(d) if g == 1]
                    dims.append(dv[0] if dv else 0)
            n_features = dims[1]
    return node_attrs, n_features


def _packed_ints(wt, v, acc):
    if wt == 2:
        i = 0
        while i < len(v):
            x, i = _varint(v, i)
            acc.append(_signed(x))
    else:
        acc.append(_signed(v))


def initializers(model_bytes):
    """[(name, dims, data_type, float values)] of graph.initializer (GraphProto.initializer = 5; TensorProto.dims = 1,
    data_type = 2, float_data = 4, name = 8, raw_data = 9; data_type 1 = FLOAT)."""
    graph = next(v for fn, _, v in fields(model_bytes) if fn == 7)
    out = []
    for fn, _, v in fields(graph):
        if fn != 5:
            continue
        name, dims, dtype, floats, raw = "", [], 0, [], b""
        for f, wt, val in fields(v):
            if f == 1:
                _packed_ints(wt, val, dims)
            elif f == 2:
                dtype = val
            elif f == 4:
                floats += list(struct.unpack(f"<{len(val) // 4}f", val)) if wt == 2 else [struct.unpack("<f", val)[0]]
            elif f == 8:
                name = val.decode()
            elif f == 9:
                raw = val
        if not floats and raw and dtype == 1:
            floats = list(struct.unpack(f"<{len(raw) // 4}f", raw))
        out.append((name, dims, dtype, floats))
    return out


def nodes(model_bytes):
    """[(name, op_type, outputs)] of graph.node (NodeProto.output = 2, name = 3, op_type = 4)."""
    graph = next(v for fn, _, v in fields(model_bytes) if fn == 7)
    out = []
    for fn, _, v in fields(graph):
        if fn == 1:
            name, op, outs = "", "", []
            for f, _, val in fields(v):
                if f == 2:
                    outs.append(val.decode())
                elif f == 3:
                    name = val.decode()
                elif f == 4:
                    op = val.decode()
            out.append((name, op, outs))
    return out


def input_features(model_bytes):
    """dim_value of axis 1 of graph.input[0]."""
    graph = next(v for fn, _, v in fields(model_bytes) if fn == 7)
    first = next(v for fn, _, v in fields(graph) if fn == 11)
    ty = next(val for f, _, val in fields(first) if f == 2)
    tensor = next(val for f, _, val in fields(ty) if f == 1)
    shape = next(val for f, _, val in fields(tensor) if f == 2)
    dims = []
    for f, _, d in fields(shape):
        if f == 1:
            dv = [val for g, _, val in fields(d) if g == 1]
            dims.append(dv[0] if dv else 0)
    return dims[1]


def producer_name(model_bytes) -> str:
    """ModelProto.producer_name = 2."""
    for fn, wt, v in fields(model_bytes):
        if fn == 2 and wt == 2:
            return v.decode()
    return ""
