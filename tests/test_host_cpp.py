"""C++ host layer (host/): BackendB200 + the batch API bodies, compiled against the reference's own headers.
CPU: the library exports the reference's batch-API and backend symbols.  GPU: host/_build/host_parity drives the real
reference (keygen, encrypt, CPU results) and compares every output of the host layer with it, bit for bit."""
import json
import os
import subprocess

import pytest

from conftest import ROOT

BUILD = os.path.join(ROOT, "host", "_build")
LIB = os.path.join(BUILD, "libluxb200_host.so")
BIN = os.path.join(BUILD, "host_parity")

needs_build = pytest.mark.skipif(not os.path.exists(LIB), reason="host/_build not built (needs /root/reference headers: make -C host)")


@needs_build
def test_host_library_exports_reference_api():
    out = subprocess.run(["nm", "-DC", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    for sym in ["lux::fhe::EvalBinGateBatch(", "lux::fhe::BootstrapBatch(", "lux::fhe::EvalFuncBatch(",
                "lux::fhe::EvalFuncMultiOutputBatch(", "lux::fhe::EvalCMUXBatch(", "lux::fhe::KeySwitchBatch(",
                "lux::fhe::ModSwitchBatch(", "lux::fhe::BootstrapBatchAsync(", "lux::fhe::BatchDAG::Execute(",
                "lux::fhe::backend::BackendB200::BlindRotateBatch(", "lux::fhe::backend::BackendB200::ExternalProductBatch(",
                "lux::fhe::backend::BackendB200::KeySwitchBatch(", "lux::fhe::backend::BackendB200::ModSwitchBatch(",
                "lux::fhe::backend::BackendB200::PackBootstrappingKey(", "lux::fhe::backend::RegisterB200("]:
        assert sym in out, sym


@needs_build
def test_host_layer_refuses_without_gpu():
    from conftest import has_gpu
    if has_gpu():
        pytest.skip("GPU present")
    r = subprocess.run([BIN, "TOY", "2", "0"], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


@pytest.mark.gpu
@needs_build
@pytest.mark.parametrize("pset,cnt,pairs", [("TOY", 12, 2), ("TOY_AP", 12, 1), ("TOY_LMKCDEY", 12, 1), ("STD128", 9, 2),
                                            ("STD128_LMKCDEY", 6, 1), ("STD128_3", 6, 1), ("STD128_4_LMKCDEY", 6, 1)])
def test_host_layer_matches_reference(pset, cnt, pairs):
    r = subprocess.run([BIN, pset, str(cnt), str(pairs)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    assert res["failed_checks"] == 0
    for k in ["BootstrapBatch", "EvalCMUXBatch", "BatchDAG_levels", "GreaterThan8_circuit", "Backend_BlindRotateBatch",
              "Backend_ExternalProductBatch", "Backend_KeySwitchBatch", "Backend_ModSwitchBatch", "Backend_PackUnpackBSK"]:
        assert res[k] is True, k


CAPI = os.path.join(BUILD, "c_api_parity")


@needs_build
def test_c_abi_batch_symbols_and_refusal():
    """include/lux/fhe/c_api_batch.h: every declared symbol is exported next to the reference's own lux_fhe_* calls, and
    without a CUDA device lux_fhe_gpu_enable fails (the batch calls never run the CPU loop)."""
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    for sym in ["lux_fhe_gpu_enable", "lux_fhe_mux_batch", "lux_fhe_bootstrap_batch", "lux_fhe_and", "lux_fhe_context_new"] + \
               [f"lux_fhe_{g}_batch" for g in ("and", "or", "xor", "nand", "nor", "xnor")]:
        assert f" T {sym}\n" in out, sym
    from conftest import has_gpu
    if not has_gpu():
        r = subprocess.run([CAPI, "toy", "2"], capture_output=True, text=True, timeout=600)
        assert r.returncode == 3 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
@needs_build
@pytest.mark.parametrize("pset,cnt", [("toy", 8), ("std128", 8)])
def test_c_abi_batch_matches_single_calls(pset, cnt):
    """plain-C program over the FFI: lux_fhe_<gate>_batch / mux_batch / bootstrap_batch (B200) vs the reference's
    lux_fhe_<gate> / mux / bootstrap (CPU), compared through lux_fhe_ciphertext_marshal bytes."""
    r = subprocess.run([CAPI, pset, str(cnt)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert json.loads(r.stdout.strip().splitlines()[-1])["failed_checks"] == 0


PACKED = os.path.join(BUILD, "packed_parity")


@needs_build
@pytest.mark.parametrize("pset", ["TOY", "TOY_AP", "TOY_LMKCDEY"])
def test_packed_key_formats_roundtrip_cpu(pset):
    """backend/packed.h key formats (host/packed_b200.cpp; the reference leaves them empty, packed.cpp:284-331): reference
    keygen -> Pack -> Unpack gives identical keys, malformed bytes are refused, and the reference's CPU gate on the
    unpacked keys equals the gate on the original keys.  No GPU involved."""
    r = subprocess.run([PACKED, pset, "4", "0"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert json.loads(r.stdout.strip().splitlines()[-1])["failed_checks"] == 0


@pytest.mark.gpu
@needs_build
@pytest.mark.parametrize("pset,cnt", [("TOY", 16), ("TOY_AP", 8), ("TOY_LMKCDEY", 8), ("STD128", 8), ("STD128_LMKCDEY", 6)])
def test_packed_keys_load_on_b200(pset, cnt):
    """a context that never held keys: BackendB200::LoadPacked(bytes) then EvalBinGateBatch == reference CPU gates."""
    r = subprocess.run([PACKED, pset, str(cnt), "1"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "B200 gates on LoadPacked keys == reference" in r.stdout
    assert json.loads(r.stdout.strip().splitlines()[-1])["failed_checks"] == 0
