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


NEW_CHECKS = ["EvalBinGateBatch_extended_inputs", "EvalBinGateBatch_same_ct_refused", "EvalBinGateBatch_ptmodulus",
              "Async_overlapping_futures", "Rekey_follows_context", "Second_context_rebinds"]


def _run_parity(args, timeout=1500):
    r = subprocess.run([BIN] + [str(a) for a in args], capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    assert res["failed_checks"] == 0
    keys = ["BootstrapBatch", "EvalCMUXBatch", "BatchDAG_levels", "Backend_BlindRotateBatch", "Backend_ExternalProductBatch",
            "Backend_KeySwitchBatch", "Backend_ModSwitchBatch", "Backend_PackUnpackBSK"] + NEW_CHECKS
    if int(args[2]) > 0:
        keys.append("GreaterThan8_circuit")
    for k in keys:
        assert res[k] is True, k
    return res, r.stdout


@pytest.mark.gpu
@needs_build
@pytest.mark.parametrize("pset,cnt,pairs", [("TOY", 12, 2), ("TOY_AP", 12, 1), ("TOY_LMKCDEY", 12, 1), ("STD128", 9, 2),
                                            ("STD128_LMKCDEY", 6, 1), ("STD128_3", 6, 1), ("STD128_4_LMKCDEY", 6, 1),
                                            ("STD128_AP", 4, 1),    # STD128_AP: the reference's own 1 GB refresh key
                                            ("STD256", 2, 0), ("STD256_3", 2, 0),   # N = 2048: four warps per ciphertext,
                                            ("STD256_LMKCDEY", 2, 0), ("STD192", 2, 0)])   # and the shared-memory kernel (64-bit Q)
def test_host_layer_matches_reference(pset, cnt, pairs):
    _run_parity([pset, cnt, pairs])


@pytest.mark.gpu
@needs_build
@pytest.mark.parametrize("pset,cnt,pairs", [("TOY", 12, 2), ("TOY_AP", 8, 1), ("TOY_LMKCDEY", 8, 1), ("STD128", 6, 1),
                                            ("STD128_LMKCDEY", 6, 1)])
def test_device_keygen_serves_the_reference(pset, cnt, pairs):
    """BackendB200::BTKeyGen: the refresh and switching keys are generated on the B200 and handed to the reference's context
    (BTKeyLoad).  The whole parity program then runs on them: the reference's CPU evaluator and the device agree bit for bit
    on device-generated keys, and results decrypt correctly under the reference's secret key."""
    _run_parity([pset, cnt, pairs, "--gpu-keygen"])


@pytest.mark.gpu
@needs_build
def test_dcrtpoly_switchformat_seam():
    """host/dcrt_b200: DCRTPoly::SwitchFormat (dcrtpoly-impl.h:1953-1959) at N = 2^16, 32 limbs of 60 bits, every limb of
    the reference's own DCRTPoly objects, forward and inverse."""
    res, _ = _run_parity(["TOY", 6, 0, "--dcrt"])
    assert res["DCRT_SwitchFormat_forward"] is True and res["DCRT_SwitchFormat_inverse"] is True


@pytest.mark.gpu
@needs_build
def test_evalbingatebatch_e2e_cpp():
    """the call a reference user makes — lux::fhe::EvalBinGateBatch(cc, AND, vector<LWECiphertext> ...) — on 2^15 gates:
    every output decrypts to the truth table, the first 64 are bit-identical to cc.EvalBinGate, throughput is printed."""
    res, out = _run_parity(["STD128", 6, 0, "--e2e", 32768])
    assert res["e2e_cpp_correct"] is True
    line = [ln for ln in out.splitlines() if ln.startswith('{"e2e_cpp"')][-1]
    assert json.loads(line)["e2e_cpp"]["bootstraps_per_s"] > 1000


@pytest.mark.gpu
@needs_build
def test_multi_device_backend():
    """one BackendB200 over 2 GPUs (keys copied peer-to-peer with luxb200_broadcast_keys, batches split over the devices):
    the same parity program, every check, against the live reference.  Needs a 2-GPU box (gpurun --gpus 2)."""
    from fhe_b200 import lib
    if lib.load().luxb200_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    res, _ = _run_parity(["STD128", 9, 2, "--devices", 2])
    assert res["devices"] == 2
    res, _ = _run_parity(["TOY_LMKCDEY", 12, 1, "--devices", 2])
    assert res["devices"] == 2


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


LARGEPREC = os.path.join(BUILD, "largeprec_parity")


@pytest.mark.gpu
@needs_build
@pytest.mark.parametrize("logQ,time_opt", [(17, 1), (27, 1), (23, 0)])
def test_largeprec_batch_layer_matches_reference(logQ, time_opt):
    """host/largeprec_b200: EvalFloorBatch / EvalSignBatch / EvalDecompBatch on the reference's own large-precision context
    (keys from BTKeyGen — three key sets with timeOptimization — and Encrypt at Q = 2^logQ) vs cc.EvalFloor / EvalSign /
    EvalDecomp (binfhecontext.cpp:346-370), ciphertext for ciphertext."""
    r = subprocess.run([LARGEPREC, "TOY", str(logQ), str(time_opt), "8"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    assert res["failed_checks"] == 0 and res["key_sets"] == (3 if time_opt else 1)
    for k in ["EvalFloorBatch", "EvalSignBatch", "EvalSignBatch_schemeSwitch", "EvalDecompBatch", "EvalSignBatch_refuses_small"]:
        assert res[k] is True, k


@needs_build
def test_largeprec_layer_refuses_without_gpu():
    from conftest import has_gpu
    if has_gpu():
        pytest.skip("GPU present")
    r = subprocess.run([LARGEPREC, "TOY", "13", "0", "2"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout
