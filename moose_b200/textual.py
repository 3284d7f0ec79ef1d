"""Reader for Moose's textual computation format, host dialect.

Grammar restated from moose/src/textual/parsing.rs (nom parser) and the printer in moose/src/textual/mod.rs;
op/placement model from moose/src/computation.rs:1656-1661.  One operation per line:

    name = Kind{attr = value, ...}: (InTy, ...) -> OutTy (input, ...) @Host(role)
    name = AddN: [HostRing128Tensor] -> HostRing128Tensor (a, b, c) @Host(role)

Only what lowered (host-dialect) graphs use is interpreted: the executors only ever see such graphs
(moose/src/compilation/lowering.rs:4-6).  Logical-dialect lines (Tensor<Fixed128(24, 40)>, @Replicated(..))
parse too -- the structure is the same -- but are not executable here.
"""
from __future__ import annotations

import ast
import json
import re
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple


@dataclass
class Operation:
    name: str
    kind: str
    attrs: Dict[str, str]
    input_types: List[str]
    variadic: bool
    output_type: str
    inputs: List[str]
    placement_kind: str
    owners: List[str]

    def to_json(self):
        return [self.name, self.kind, self.attrs, self.input_types, int(self.variadic), self.output_type, self.inputs,
                self.placement_kind, self.owners]

    def to_text(self) -> str:
        """One line of the textual format (the printer's layout, moose/src/textual/mod.rs)."""
        attrs = "{" + ", ".join(f"{k} = {v}" for k, v in self.attrs.items()) + "}" if self.attrs else ""
        types = ", ".join(self.input_types)
        sig = f"[{types}]" if self.variadic else f"({types})"
        return (f"{self.name} = {self.kind}{attrs}: {sig} -> {self.output_type} ({', '.join(self.inputs)}) "
                f"@{self.placement_kind}({', '.join(self.owners)})")

    @staticmethod
    def from_json(row) -> "Operation":
        return Operation(row[0], row[1], row[2], row[3], bool(row[4]), row[5], row[6], row[7], row[8])


class ParseError(ValueError):
    pass


_OPEN, _CLOSE = "([{<", ")]}>"


def _split_top(s: str, sep: str = ",") -> List[str]:
    """Split at top-level separators (outside brackets and double quotes)."""
    out, depth, cur, in_str = [], 0, [], False
    i = 0
    while i < len(s):
        ch = s[i]
        if in_str:
            cur.append(ch)
            if ch == "\\" and i + 1 < len(s):
                cur.append(s[i + 1])
                i += 1
            elif ch == '"':
                in_str = False
        elif ch == '"':
            in_str = True
            cur.append(ch)
        elif ch in _OPEN:
            depth += 1
            cur.append(ch)
        elif ch in _CLOSE and not (ch == ">" and i > 0 and s[i - 1] == "-"):
            depth -= 1
            cur.append(ch)
        elif ch == sep and depth == 0:
            out.append("".join(cur).strip())
            cur = []
        else:
            cur.append(ch)
        i += 1
    last = "".join(cur).strip()
    if last or out:
        out.append(last)
    return [p for p in out if p != ""]


def _matching_open(s: str, close_idx: int) -> int:
    """Index of the bracket matching s[close_idx] (scanning backwards, quotes respected)."""
    depth, in_str = 0, False
    i = close_idx
    while i >= 0:
        ch = s[i]
        if in_str:
            if ch == '"' and (i == 0 or s[i - 1] != "\\"):
                in_str = False
        elif ch == '"':
            in_str = True
        elif ch in _CLOSE and not (ch == ">" and i > 0 and s[i - 1] == "-"):
            depth += 1
        elif ch in _OPEN:
            depth -= 1
            if depth == 0:
                return i
        i -= 1
    raise ParseError(f"unbalanced brackets in {s!r}")


def _matching_close(s: str, open_idx: int) -> int:
    depth, in_str = 0, False
    for i in range(open_idx, len(s)):
        ch = s[i]
        if in_str:
            if ch == '"' and s[i - 1] != "\\":
                in_str = False
        elif ch == '"':
            in_str = True
        elif ch in _OPEN:
            depth += 1
        elif ch in _CLOSE and not (ch == ">" and s[i - 1] == "-"):
            depth -= 1
            if depth == 0:
                return i
    raise ParseError(f"unbalanced brackets in {s!r}")


def parse_line(line: str) -> Operation:
    line = line.strip()
    m = re.match(r"^([A-Za-z_][A-Za-z0-9_\-]*)\s*=\s*([A-Za-z][A-Za-z0-9]*)", line)
    if not m:
        raise ParseError(f"cannot parse operation: {line[:80]!r}")
    name, kind = m.group(1), m.group(2)
    rest = line[m.end():].lstrip()
    attrs: Dict[str, str] = {}
    if rest.startswith("{"):
        end = _matching_close(rest, 0)
        for part in _split_top(rest[1:end]):
            if "=" not in part:
                raise ParseError(f"bad attribute {part!r} in {name}")
            k, v = part.split("=", 1)
            attrs[k.strip()] = v.strip()
        rest = rest[end + 1:].lstrip()
    if not rest.startswith(":"):
        raise ParseError(f"expected ':' after operator in {name}")
    rest = rest[1:].strip()
    at = rest.rfind(" @")
    if at < 0:
        raise ParseError(f"missing placement in {name}")
    plc = rest[at + 2:].strip()
    body = rest[:at].rstrip()
    pm = re.match(r"^([A-Za-z0-9]+)\((.*)\)$", plc)
    if not pm:
        raise ParseError(f"bad placement {plc!r} in {name}")
    placement_kind, owners = pm.group(1), [o.strip() for o in _split_top(pm.group(2))]
    # signature first, then the OPTIONAL argument list (parsing.rs:176 `opt(argument_list)`: a nullary operation may
    # be written without "()"); a type is alphanumeric with an optional <...> (parsing.rs:497-507), never "(...)"
    variadic = body.startswith("[")
    if not (body.startswith("(") or variadic):
        raise ParseError(f"bad signature {body!r} in {name}")
    end = _matching_close(body, 0)
    input_types = _split_top(body[1:end])
    after = body[end + 1:].strip()
    if not after.startswith("->"):
        raise ParseError(f"missing '->' in signature of {name}")
    after = after[2:].strip()
    tm = re.match(r"^[A-Za-z0-9]+(<[^>]*>)?", after)
    if not tm:
        raise ParseError(f"missing output type in {name}")
    output_type = tm.group(0)
    tail = after[tm.end():].strip()
    if not tail:
        inputs = []
    elif tail.startswith("(") and _matching_close(tail, 0) == len(tail) - 1:
        inputs = _split_top(tail[1:-1])
    else:
        raise ParseError(f"bad input list {tail!r} in {name}")
    return Operation(name, kind, attrs, input_types, variadic, output_type, inputs, placement_kind, owners)


def parse_computation(text: str) -> List[Operation]:
    ops = []
    for raw in text.splitlines():
        line = raw.strip()
        if not line or line.startswith("//") or line.startswith("#"):
            continue
        ops.append(parse_line(line))
    return ops


def dumps_text(ops: List[Operation]) -> str:
    return "\n".join(op.to_text() for op in ops) + "\n"


def dumps_json(ops: List[Operation]) -> str:
    return json.dumps([op.to_json() for op in ops], separators=(",", ":"))


def loads_json(s: str) -> List[Operation]:
    return [Operation.from_json(r) for r in json.loads(s)]


# ---- attribute / constant value decoding -------------------------------------------------------

def attr_int(op: Operation, key: str) -> int:
    return int(op.attrs[key])


def attr_int_list(op: Operation, key: str) -> List[int]:
    """`axis = [0, 2]` (ExpandDims, parsing.rs:327-331)."""
    return [int(v) for v in ast.literal_eval(op.attrs[key])]


def attr_slice(op: Operation, key: str) -> List[Tuple[int, Optional[int], Optional[int]]]:
    """`slice = {start = 1, end = 10, step = -1}`: the textual form carries ONE SliceInfoElem
    (slice_info_literal, parsing.rs:902-913), end and step optional."""
    body = op.attrs[key].strip()
    if not (body.startswith("{") and body.endswith("}")):
        raise ParseError(f"bad slice literal {body!r} in {op.name}")
    fields = {}
    for part in _split_top(body[1:-1]):
        k, v = part.split("=", 1)
        fields[k.strip()] = int(v)
    if "start" not in fields:
        raise ParseError(f"slice literal without start in {op.name}")
    return [(fields["start"], fields.get("end"), fields.get("step"))]


def attr_str(op: Operation, key: str) -> str:
    v = op.attrs[key]
    return ast.literal_eval(v) if v.startswith('"') else v


def attr_hex16(op: Operation, key: str) -> bytes:
    """sync_key / rendezvous_key: 32 hex digits, or a byte list like [1, 2, 3] zero-padded to 16
    (SyncKey::try_from, moose/src/host/prim.rs:98-111)."""
    v = op.attrs[key].strip()
    if v.startswith("["):
        b = bytes(ast.literal_eval(v))
        return b + bytes(16 - len(b))
    b = bytes.fromhex(v)
    return b + bytes(16 - len(b))


def parse_constant(value: str) -> Tuple[str, object]:
    """`Ty(payload)` -> (Ty, python value).  Tensors become nested lists, keys / seeds bytes."""
    m = re.match(r"^([A-Za-z0-9]+)\((.*)\)$", value.strip(), re.S)
    if not m:
        raise ParseError(f"bad constant {value!r}")
    ty, payload = m.group(1), m.group(2).strip()
    if ty in ("HostPrfKey", "HostSeed", "PrfKey", "Seed"):
        return ty, bytes.fromhex(payload)
    if ty in ("HostString", "String"):
        return ty, ast.literal_eval(payload)
    if ty in ("HostShape", "Shape"):
        return ty, tuple(ast.literal_eval(payload))
    return ty, ast.literal_eval(payload)
