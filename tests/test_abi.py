"""CPU-only checks of the C-ABI boundary: the library loads, exports every symbol that
include/moose_b200.h declares, and its host-side (CPU) control-value code -- BLAKE3 seed
derivation, session ids, the AES key schedule -- agrees with the oracle and with the
reference's golden vectors.  No compute calls: there is no GPU here."""
import ctypes
import os
import re

import pytest

import moose_b200
from oracle import moose_oracle as mo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "moose_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = moose_b200.lib()
    syms = declared_symbols()
    assert len(syms) > 60
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    # the python mirror binds exactly the declared ABI
    assert sorted(moose_b200.ABI.keys()) == syms


def test_no_cpu_fallback_without_device():
    from _util import HAS_GPU
    if HAS_GPU:
        pytest.skip("GPU present")
    with pytest.raises(moose_b200.MooseB200Error) as e:
        moose_b200.Session(device=0)
    assert e.value.code == -6  # MB_ERR_NO_DEVICE


def test_host_side_derive_seed_matches_reference_vector():
    lib = moose_b200.lib()
    sid = ctypes.create_string_buffer(16)
    lib.mb_selftest_session_id(b"foobar", sid)
    assert sid.raw == mo.session_id_bytes("foobar")
    out = ctypes.create_string_buffer(16)
    lib.mb_selftest_derive_seed(sid.raw, mo.sync_key_from([1, 2, 3]), bytes(16), out)
    # moose/src/execution/mod.rs:172-199
    assert list(out.raw) == [79, 203, 243, 208, 77, 199, 116, 216, 2, 206, 173, 36, 20, 204, 200, 146]


def test_host_side_blake3_against_independent_implementation():
    import numpy as np
    lib = moose_b200.lib()
    rng = np.random.default_rng(7)
    for n in (0, 1, 11, 63, 64, 65, 200, 1024):
        s = bytes(rng.integers(1, 255, n, dtype=np.uint8))  # no NUL: passed as a C string
        out = ctypes.create_string_buffer(16)
        lib.mb_selftest_session_id(s, out)
        assert out.raw == mo.session_id_bytes(s.decode("latin1")) or out.raw == __import__("blake3").blake3(s).digest(16)
    for _ in range(20):
        sid, sync, key = (bytes(rng.integers(0, 256, 16, dtype=np.uint8)) for _ in range(3))
        out = ctypes.create_string_buffer(16)
        lib.mb_selftest_derive_seed(sid, sync, key, out)
        assert out.raw == mo.derive_seed(sid, sync, key)


def test_host_side_aes_key_schedule():
    lib = moose_b200.lib()
    out = ctypes.create_string_buffer(16)
    lib.mb_selftest_aes_block(bytes(range(16)), bytes.fromhex("00112233445566778899aabbccddeeff"), out)
    assert out.raw.hex() == "69c4e0d86a7b0430d8cdb78070b4c55a"  # FIPS-197 C.1
    lib.mb_selftest_aes_block(bytes(16), bytes(16), out)
    assert out.raw == mo.aes_ctr(bytes(16), 0, 1)


def test_generated_rust_ffi_is_in_sync():
    """integration/rust/b200_ffi.rs (the extern block for the Rust shim) is generated from the
    header: regenerating must reproduce the committed file and cover every ABI entry."""
    import re
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = os.path.join(root, "integration", "rust", "b200_ffi.rs")
    before = open(path).read()
    subprocess.run([sys.executable, os.path.join(root, "tools", "gen_rust_ffi.py")], check=True, capture_output=True)
    assert open(path).read() == before
    import moose_b200 as mb
    assert set(re.findall(r"pub fn (mb_\w+)\(", before)) == set(mb.ABI)


def test_rust_shim_source_calls_only_declared_entry_points():
    """integration/rust/host_b200.rs (the kernel bodies a Moose maintainer adds; source only, no Rust toolchain here)
    must bind entry points that exist in the generated extern block and therefore in the header."""
    src = open(os.path.join(ROOT, "integration", "rust", "host_b200.rs")).read()
    ffi = open(os.path.join(ROOT, "integration", "rust", "b200_ffi.rs")).read()
    code = re.sub(r"//.*", "", src)                                   # identifiers in code, macro arguments included
    used = set(re.findall(r"\b(mb_[a-z0-9_]+)\b", code))
    declared = set(re.findall(r"pub fn (mb_[a-z0-9_]+)", ffi))
    assert len(used) > 25
    assert not (used - declared), sorted(used - declared)
    assert declared == set(declared_symbols())      # the generated block tracks the header
