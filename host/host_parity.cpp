// host_parity.cpp — drop-in check of the C++ host layer: the reference's own BinFHEContext generates parameters,
// keys and ciphertexts; the batch API (host/batch_b200.cpp) and the backend seam (host/backend_b200.cpp) run on the
// B200; every output is compared coefficient by coefficient with what the reference's CPU code returns for the same
// inputs (cc.EvalBinGate / Bootstrap / EvalFunc / EvalAcc / KeySwitch / ModSwitch).  TEST INFRASTRUCTURE: links the
// reference library built by oracle/Makefile.
// usage: host_parity <TOY|STD128|STD128_AP|STD128_LMKCDEY|...> [count] [pairs] [--devices G] [--e2e GATES] [--dcrt]
//   --devices G  one BackendB200 over GPUs 0..G-1 (keys copied peer-to-peer, every batch split over the devices)
//   --e2e GATES  also time lux::fhe::EvalBinGateBatch(cc, AND, ...) on GATES LWECiphertext pairs (the call a reference user
//                makes: reference containers in, reference containers out) and print e2e_cpp
//   --dcrt       also check host/dcrt_b200 against DCRTPoly::SwitchFormat at N = 2^16, 32 limbs of 60 bits
#include <cstdio>
#include <execinfo.h>
#include <csignal>
#include <unistd.h>
#include <cstdlib>
#include <functional>
#include <random>
#include <string>

#include <chrono>
#include <cstring>

#include "backend_b200.h"
#include "dcrt_b200.h"
#include "batch/binfhe-batch.h"
#include "rgsw-acc-cggi.h"
#include "rgsw-acc-dm.h"
#include "rgsw-acc-lmkcdey.h"

using namespace lux::fhe;

static int g_fail = 0;
static std::string g_json;

static bool same_ct(const LWECiphertext& x, const LWECiphertext& y) {
    return x && y && x->GetLength() == y->GetLength() && x->GetModulus() == y->GetModulus() && x->GetB() == y->GetB() &&
           x->GetA() == y->GetA();
}
static void report(const char* name, size_t n, size_t bad) {
    std::printf("%-28s %4zu compared, %zu mismatches\n", name, n, bad);
    if (bad) ++g_fail;
    g_json += std::string(g_json.empty() ? "" : ", ") + "\"" + name + "\": " + (bad ? "false" : "true");
}
static size_t count_diff(const std::vector<LWECiphertext>& a, const std::vector<LWECiphertext>& b) {
    if (a.size() != b.size()) return std::max(a.size(), b.size());
    size_t bad = 0;
    for (size_t i = 0; i < a.size(); ++i) bad += !same_ct(a[i], b[i]);
    return bad;
}

// GreaterThan8 of go/tfhe/compare.go:106-139 as levelled batches over `pairs` independent comparisons: the 3 gates per
// bit that do not depend on the running result go out as one batch of 24*pairs gates, then 3 dependent levels per bit.
using GateBatch = std::function<std::vector<LWECiphertext>(BINGATE, const std::vector<LWECiphertext>&, const std::vector<LWECiphertext>&)>;
static std::vector<LWECiphertext> greater_than8(BinFHEContext& cc, const std::vector<std::vector<LWECiphertext>>& a,
                                                const std::vector<std::vector<LWECiphertext>>& b,
                                                const std::vector<LWECiphertext>& init, const GateBatch& gate) {
    const size_t P = a.size();
    std::vector<LWECiphertext> x1, y1, x2, y2, x3, y3;
    for (size_t p = 0; p < P; ++p)
        for (int i = 0; i < 8; ++i) {
            x1.push_back(a[p][i]); y1.push_back(cc.EvalNOT(b[p][i]));   // aGreater = a AND NOT b
            x2.push_back(cc.EvalNOT(a[p][i])); y2.push_back(b[p][i]);   // bGreater = NOT a AND b
            x3.push_back(a[p][i]); y3.push_back(b[p][i]);               // equal = a XNOR b
        }
    auto aG = gate(AND, x1, y1), bG = gate(AND, x2, y2), eq = gate(XNOR, x3, y3);
    std::vector<LWECiphertext> result = init;
    for (int i = 7; i >= 0; --i) {
        std::vector<LWECiphertext> e(P), g(P), nb(P);
        for (size_t p = 0; p < P; ++p) { e[p] = eq[p * 8 + i]; g[p] = aG[p * 8 + i]; nb[p] = cc.EvalNOT(bG[p * 8 + i]); }
        result = gate(AND, e, result);
        result = gate(OR, g, result);
        result = gate(AND, result, nb);
    }
    return result;
}
static int greater_than8_plain(unsigned a, unsigned b) {   // the same circuit on plaintext bits
    int r = 0;
    for (int i = 7; i >= 0; --i) {
        const int ai = (a >> i) & 1, bi = (b >> i) & 1;
        r = (ai & !bi) | ((ai == bi) & r);
        r = r & !((!ai) & bi);
    }
    return r;
}

static double wall_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// DCRTPolyImpl::SwitchFormat (dcrtpoly-impl.h:1953-1959) on the CPU vs host/dcrt_b200 (one luxb200_ntt_batch call), every limb
static void dcrt_check(uint32_t N, uint32_t L, uint32_t bits, size_t count) {
    auto params = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, L, bits);
    std::mt19937_64 rng(99);
    std::vector<DCRTPoly> gpu, cpu;
    for (size_t i = 0; i < count; ++i) {
        DCRTPoly p(params, Format::COEFFICIENT, true);
        for (uint32_t l = 0; l < L; ++l) {
            NativePoly& limb = p.GetAllElements()[l];
            const uint64_t q = limb.GetModulus().ConvertToInt<uint64_t>();
            for (uint32_t k = 0; k < N; ++k) limb[k] = NativeInteger(rng() % q);
            if (i == 0 && l == 0) for (uint32_t k = 0; k < N; ++k) limb[k] = NativeInteger(q - 1);
        }
        gpu.push_back(p);
        cpu.push_back(std::move(p));
    }
    backend::DCRTNttB200 ntt(0);
    std::vector<DCRTPoly*> ptrs;
    for (auto& p : gpu) ptrs.push_back(&p);
    const double t0 = wall_ms();
    ntt.SwitchFormat(ptrs);
    const double t1 = wall_ms();
    for (auto& p : cpu) p.SwitchFormat();
    const double t2 = wall_ms();
    size_t bad = 0;
    for (size_t i = 0; i < count; ++i) bad += !(gpu[i] == cpu[i]) || gpu[i].GetFormat() != Format::EVALUATION;
    report("DCRT_SwitchFormat_forward", count * L, bad);
    ntt.SwitchFormat(ptrs);
    for (auto& p : cpu) p.SwitchFormat();
    bad = 0;
    for (size_t i = 0; i < count; ++i) bad += !(gpu[i] == cpu[i]) || gpu[i].GetFormat() != Format::COEFFICIENT;
    report("DCRT_SwitchFormat_inverse", count * L, bad);
    std::printf("DCRT N=%u L=%u x%zu: B200 seam %.1f ms (host packing + PCIe included), reference CPU %.1f ms\n", N, L, count, t1 - t0,
                t2 - t1);
}

// a crash of the acceptance program names its frames on stderr instead of dying silently (stdout is a pipe under pytest)
static void on_fatal_signal(int sig) {
    void* bt[64];
    const int nbt = backtrace(bt, 64);
    std::fprintf(stderr, "host_parity: fatal signal %d\n", sig);
    backtrace_symbols_fd(bt, nbt, 2);
    _exit(128 + sig);
}

int main(int argc, char** argv) {
    std::signal(SIGSEGV, on_fatal_signal);
    std::signal(SIGABRT, on_fatal_signal);
    std::signal(SIGBUS, on_fatal_signal);
    std::setvbuf(stdout, nullptr, _IOLBF, 0);
    std::vector<std::string> pos;
    size_t n_dev = 1, e2e_gates = 0;
    bool dcrt = false, gpu_keygen = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--devices" && i + 1 < argc) n_dev = std::strtoul(argv[++i], nullptr, 10);
        else if (a == "--e2e" && i + 1 < argc) e2e_gates = std::strtoul(argv[++i], nullptr, 10);
        else if (a == "--dcrt") dcrt = true;
        else if (a == "--gpu-keygen") gpu_keygen = true;   // BTKeyGen on the device instead of the reference's
        else pos.push_back(a);
    }
    const std::string set = pos.size() > 0 ? pos[0] : "STD128";
    const size_t cnt = pos.size() > 1 ? std::strtoul(pos[1].c_str(), nullptr, 10) : 8;
    const size_t pairs = pos.size() > 2 ? std::strtoul(pos[2].c_str(), nullptr, 10) : 2;
    BINFHE_PARAMSET ps = STD128;
    BINFHE_METHOD method = GINX;
    if (set == "TOY") ps = TOY;
    else if (set == "STD128_AP") { ps = STD128_AP; method = AP; }
    else if (set == "STD128_LMKCDEY") { ps = STD128_LMKCDEY; method = LMKCDEY; }
    else if (set == "TOY_AP") { ps = TOY; method = AP; }
    else if (set == "TOY_LMKCDEY") { ps = TOY; method = LMKCDEY; }
    else if (set == "MEDIUM") ps = MEDIUM;
    else if (set == "STD128_3") ps = STD128_3;          // baseG = 128: digitsG - 1 = 3
    else if (set == "STD128Q") ps = STD128Q;
    else if (set == "STD128_4_LMKCDEY") { ps = STD128_4_LMKCDEY; method = LMKCDEY; }
    else if (set == "STD256") ps = STD256;              // N = 2048, 29-bit Q, 2 digits: the four-warp kernel (br_quad.cuh)
    else if (set == "STD256_3") ps = STD256_3;          // 3 digits
    else if (set == "STD256_LMKCDEY") { ps = STD256_LMKCDEY; method = LMKCDEY; }   // 30-bit Q: the shared-memory kernel
    else if (set == "STD192") ps = STD192;              // 37-bit Q: 64-bit residues

    BinFHEContext cc;
    cc.GenerateBinFHEContext(ps, method);
    auto sk = cc.KeyGen();
    if (!gpu_keygen) cc.BTKeyGen(sk);
    std::vector<int> devs;
    for (size_t d = 0; d < n_dev; ++d) devs.push_back(static_cast<int>(d));
    auto* be = backend::RegisterB200(cc, devs);
    if (gpu_keygen) {
        // the refresh and switching keys are drawn on the B200 and loaded into cc (BTKeyLoad): every check below then runs the
        // reference's CPU evaluator on device-generated keys against the device itself, and decrypts with the reference's sk
        const double t0 = wall_ms();
        be->BTKeyGen(cc, sk, 0x5eedULL + cnt);
        std::printf("BTKeyGen on the device + hand-over to the context: %.1f ms\n", wall_ms() - t0);
    }
    std::printf("backend: %s, %zu device(s), device memory %.1f GB, default=%s\n", be->Name().c_str(), be->NumDevices(),
                be->DeviceMemory() / 1e9, backend::BackendTypeName(backend::CurrentBackend()->Type()).c_str());

    std::mt19937 rng(12345);
    std::vector<int> m1(cnt), m2(cnt), m3(cnt);
    std::vector<LWECiphertext> c1(cnt), c2(cnt), c3(cnt);
    for (size_t i = 0; i < cnt; ++i) {
        m1[i] = rng() & 1; m2[i] = rng() & 1; m3[i] = rng() & 1;
        c1[i] = cc.Encrypt(sk, m1[i]); c2[i] = cc.Encrypt(sk, m2[i]); c3[i] = cc.Encrypt(sk, m3[i]);
    }
    // The reference fills its static NTT twiddle tables on first use, without a lock (ChineseRemainderTransformFTTNat::
    // ForwardTransformToBitReverseInPlace -> PreCompute).  cc.BTKeyGen does that as a side effect; with --gpu-keygen no
    // reference transform has run yet, and the OpenMP loops below would race on the first one (1 crash in 60 runs of
    // TOY_LMKCDEY inside the reference's SwitchFormat).  One serial gate fills the tables for this ring.
    if (cnt) (void)cc.EvalBinGate(AND, c1[0], c2[0]);

    // ---- seam 2: batch API vs the reference's per-ciphertext calls ---------------------------------------------
    for (BINGATE g : {AND, OR, NAND, NOR, XOR, XNOR}) {
        std::vector<LWECiphertext> gpu, cpu(cnt);
        auto r = EvalBinGateBatch(cc, g, c1, c2, gpu);
        if (!r.success) { std::printf("EvalBinGateBatch failed: %s\n", r.error.c_str()); return 2; }
#pragma omp parallel for
        for (size_t i = 0; i < cnt; ++i) cpu[i] = cc.EvalBinGate(g, c1[i], c2[i]);
        report(("EvalBinGateBatch_" + std::to_string(int(g))).c_str(), cnt, count_diff(gpu, cpu));
    }
    {
        std::vector<LWECiphertext> gpu, cpu(cnt);
        auto r = BootstrapBatch(cc, c1, gpu);
        if (!r.success) { std::printf("BootstrapBatch failed: %s\n", r.error.c_str()); return 2; }
#pragma omp parallel for
        for (size_t i = 0; i < cnt; ++i) cpu[i] = cc.Bootstrap(c1[i]);
        size_t bad = count_diff(gpu, cpu);
        for (size_t i = 0; i < cnt; ++i) { LWEPlaintext d; cc.Decrypt(sk, gpu[i], &d); bad += d != m1[i]; }
        report("BootstrapBatch", cnt, bad);
    }
    {
        std::vector<LWECiphertext> gpu, cpu(cnt);
        auto r = EvalCMUXBatch(cc, c1, c2, c3, gpu);
        if (!r.success) { std::printf("EvalCMUXBatch failed: %s\n", r.error.c_str()); return 2; }
#pragma omp parallel for
        for (size_t i = 0; i < cnt; ++i) cpu[i] = cc.EvalBinGate(CMUX, std::vector<LWECiphertext>{c1[i], c2[i], c3[i]});
        report("EvalCMUXBatch", cnt, count_diff(gpu, cpu));
    }
    std::vector<NativeInteger> dag_lut;   // set when EvalFunc is available for this parameter set (q <= N)
    int dag_p = 0;
    {
        // EvalFunc: p = GetMaxPlaintextSpace, f(m) = m^3 mod p  (the reference's eval-function example)
        const int p = cc.GetMaxPlaintextSpace().ConvertToInt();
        auto fp = [](NativeInteger m, NativeInteger p1) -> NativeInteger { return (m * m * m) % p1; };
        bool lut_ok = true;
        std::vector<NativeInteger> lut;
        try { lut = cc.GenerateLUTviaFunction(fp, p); } catch (...) { lut_ok = false; }
        if (lut_ok) {
            std::vector<LWECiphertext> in(cnt), gpu, cpu(cnt);
            for (size_t i = 0; i < cnt; ++i) in[i] = cc.Encrypt(sk, i % p, LARGE_DIM, p);
            std::string cpu_err;
            try { cpu[0] = cc.EvalFunc(in[0], lut); } catch (const std::exception& e) { cpu_err = e.what(); }
            if (cpu_err.empty()) {
#pragma omp parallel for
                for (size_t i = 1; i < cnt; ++i) cpu[i] = cc.EvalFunc(in[i], lut);
            }
            auto r = EvalFuncBatch(cc, in, lut, gpu);
            if (!cpu_err.empty()) {
                report("EvalFuncBatch_error_parity", 1, r.success ? 1 : 0);   // both must refuse (q > N arbitrary LUT)
            } else {
                if (!r.success) { std::printf("EvalFuncBatch failed: %s\n", r.error.c_str()); return 2; }
                report("EvalFuncBatch", cnt, count_diff(gpu, cpu));
                dag_lut = lut; dag_p = p;
            }
        }
    }

    // ---- BatchDAG: level-scheduled on the GPU vs the same DAG through the reference's CPU calls ----------------
    {
        auto build = [&](BatchDAG& d, std::vector<size_t>& outs) {
            for (size_t i = 0; i + 2 < cnt; i += 3) {
                size_t a = d.AddCiphertext(c1[i]), b = d.AddCiphertext(c2[i + 1]), c = d.AddCiphertext(c3[i + 2]);
                size_t x = d.AddBinGate(AND, a, b), y = d.AddBinGate(OR, a, c), z = d.AddBinGate(XOR, x, y);
                size_t w = d.AddCMUX(z, x, y), u = d.AddBootstrap(w);
                outs.insert(outs.end(), {x, y, z, w, u});
            }
        };
        BatchDAG dg(cc), dc(cc);
        std::vector<size_t> og, oc;
        build(dg, og); build(dc, oc);
        auto rg = dg.Execute(), rc = dc.Execute(BATCH_CPU_FORCE);
        size_t bad = (!rg.success || !rc.success) ? 1 : 0;
        for (size_t i = 0; i < og.size() && !bad; ++i) bad += !same_ct(dg.GetResult(og[i]), dc.GetResult(oc[i]));
        report("BatchDAG_levels", og.size(), bad);
    }

    // ---- BatchDAG with EvalFunc nodes: one device circuit too (luxb200_circuit_batch_ex) — same results as the reference's
    // sequential walk (batch.cpp:429-482)
    if (!dag_lut.empty()) {
        auto build = [&](BatchDAG& d, std::vector<size_t>& outs, const std::vector<LWECiphertext>& fin) {
            for (size_t i = 0; i + 1 < cnt; i += 2) {
                size_t a = d.AddCiphertext(c1[i]), b = d.AddCiphertext(c2[i + 1]), e = d.AddCiphertext(fin[i]);
                size_t x = d.AddBinGate(NAND, a, b), f = d.AddEvalFunc(e, dag_lut), y = d.AddBinGate(XNOR, x, b);
                size_t u = d.AddBootstrap(y), f2 = d.AddEvalFunc(f, dag_lut);   // a function of a function
                outs.insert(outs.end(), {x, f, y, u, f2});
            }
        };
        std::vector<LWECiphertext> fin(cnt);
        for (size_t i = 0; i < cnt; ++i) fin[i] = cc.Encrypt(sk, i % dag_p, LARGE_DIM, dag_p);
        BatchDAG dg(cc), dc(cc);
        std::vector<size_t> og, oc;
        build(dg, og, fin); build(dc, oc, fin);
        auto rg = dg.Execute(), rc = dc.Execute(BATCH_CPU_FORCE);
        size_t bad = (!rg.success || !rc.success) ? 1 : 0;
        if (bad) std::printf("BatchDAG_host_levels: %s | %s\n", rg.error.c_str(), rc.error.c_str());
        for (size_t i = 0; i < og.size() && !bad; ++i) bad += !same_ct(dg.GetResult(og[i]), dc.GetResult(oc[i]));
        report("BatchDAG_host_levels", og.size(), bad);
    }

    // ---- 8-bit comparison circuit (BASELINE config 4), levelled ------------------------------------------------
    if (pairs) {
        std::vector<std::vector<LWECiphertext>> A(pairs, std::vector<LWECiphertext>(8)), B = A;
        std::vector<unsigned> va(pairs), vb(pairs);
        std::vector<LWECiphertext> init(pairs);
        for (size_t p = 0; p < pairs; ++p) {
            va[p] = rng() & 255; vb[p] = (p & 1) ? (rng() & 255) : (va[p] ^ (1u << (rng() % 8)));
            for (int i = 0; i < 8; ++i) { A[p][i] = cc.Encrypt(sk, (va[p] >> i) & 1); B[p][i] = cc.Encrypt(sk, (vb[p] >> i) & 1); }
            init[p] = cc.Encrypt(sk, 0);
        }
        GateBatch on_gpu = [&](BINGATE g, const std::vector<LWECiphertext>& x, const std::vector<LWECiphertext>& y) {
            std::vector<LWECiphertext> o;
            auto r = EvalBinGateBatch(cc, g, x, y, o);
            if (!r.success) throw std::runtime_error(r.error);
            return o;
        };
        GateBatch on_cpu = [&](BINGATE g, const std::vector<LWECiphertext>& x, const std::vector<LWECiphertext>& y) {
            std::vector<LWECiphertext> o(x.size());
#pragma omp parallel for
            for (size_t i = 0; i < x.size(); ++i) o[i] = cc.EvalBinGate(g, x[i], y[i]);
            return o;
        };
        auto rg = greater_than8(cc, A, B, init, on_gpu), rc = greater_than8(cc, A, B, init, on_cpu);
        size_t bad = count_diff(rg, rc);
        for (size_t p = 0; p < pairs; ++p) {
            LWEPlaintext d; cc.Decrypt(sk, rg[p], &d);
            bad += d != greater_than8_plain(va[p], vb[p]);
        }
        report("GreaterThan8_circuit", pairs, bad);
    }

    // ---- seam 1: backend virtuals vs the reference's primitives -------------------------------------------------
    auto rgsw = cc.GetParams()->GetRingGSWParams();
    auto lwe  = cc.GetParams()->GetLWEParams();
    const uint32_t N = lwe->GetN(), n = lwe->Getn();
    const NativeInteger Q = lwe->GetQ();
    std::uniform_int_distribution<uint64_t> uq(0, Q.ConvertToInt() - 1);
    auto rand_rlwe = [&]() {
        std::vector<NativePoly> res(2);
        for (int c = 0; c < 2; ++c) {
            res[c] = NativePoly(rgsw->GetPolyParams(), Format::EVALUATION, true);
            for (uint32_t k = 0; k < N; ++k) res[c][k] = NativeInteger(uq(rng));
        }
        return std::make_shared<RLWECiphertextImpl>(std::move(res));
    };
    auto same_rlwe = [&](const RLWECiphertext& x, const RLWECiphertext& y) {
        return x->GetElements()[0] == y->GetElements()[0] && x->GetElements()[1] == y->GetElements()[1];
    };
    {
        const size_t m = std::min<size_t>(cnt, 4);
        std::vector<LWECiphertext> cts(c1.begin(), c1.begin() + m);
        std::vector<RLWECiphertext> gpu(m), cpu(m);
        for (size_t i = 0; i < m; ++i) { gpu[i] = rand_rlwe(); cpu[i] = std::make_shared<RLWECiphertextImpl>(*gpu[i]); }
        be->BlindRotateBatch(rgsw, cts, cc.GetRefreshKey(), gpu);
        std::shared_ptr<RingGSWAccumulator> acc;
        if (method == AP) acc = std::make_shared<RingGSWAccumulatorDM>();
        else if (method == GINX) acc = std::make_shared<RingGSWAccumulatorCGGI>();
        else acc = std::make_shared<RingGSWAccumulatorLMKCDEY>();
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) acc->EvalAcc(rgsw, cc.GetRefreshKey(), cpu[i], cts[i]->GetA());
        size_t bad = 0;
        for (size_t i = 0; i < m; ++i) bad += !same_rlwe(gpu[i], cpu[i]);
        report("Backend_BlindRotateBatch", m, bad);
    }
    {
        // ExternalProduct = one AddToAccDM step (rgsw-acc-dm.cpp:118-143) written with the reference's own primitives
        const size_t m = std::min<size_t>(cnt, 4);
        const uint32_t digitsG2 = (rgsw->GetDigitsG() - 1) << 1;
        std::vector<RingGSWEvalKey> keys(m);
        std::vector<RLWECiphertext> in(m), gpu, cpu(m);
        RingGSWAccumulatorDM dm;
        for (size_t i = 0; i < m; ++i) {
            in[i] = rand_rlwe();
            keys[i] = std::make_shared<RingGSWEvalKeyImpl>(digitsG2, 2);
            for (uint32_t r = 0; r < digitsG2; ++r)
                for (int c = 0; c < 2; ++c) {
                    NativePoly pl(rgsw->GetPolyParams(), Format::EVALUATION, true);
                    for (uint32_t k = 0; k < N; ++k) pl[k] = NativeInteger(uq(rng));
                    (*keys[i])[r][c] = std::move(pl);
                }
            std::vector<NativePoly> ct(in[i]->GetElements());
            ct[0].SetFormat(Format::COEFFICIENT); ct[1].SetFormat(Format::COEFFICIENT);
            std::vector<NativePoly> dct(digitsG2, NativePoly(rgsw->GetPolyParams(), Format::COEFFICIENT, true));
            dm.SignedDigitDecompose(rgsw, ct, dct);
            for (auto& d : dct) d.SetFormat(Format::EVALUATION);
            std::vector<NativePoly> res(2);
            for (int c = 0; c < 2; ++c) {
                res[c] = dct[0] * (*keys[i])[0][c];
                for (uint32_t l = 1; l < digitsG2; ++l) res[c] += dct[l] * (*keys[i])[l][c];
            }
            cpu[i] = std::make_shared<RLWECiphertextImpl>(std::move(res));
        }
        be->ExternalProductBatch(rgsw, keys, in, gpu);
        size_t bad = 0;
        for (size_t i = 0; i < m; ++i) bad += !same_rlwe(gpu[i], cpu[i]);
        report("Backend_ExternalProductBatch", m, bad);
    }
    {
        // KeySwitch on (N, qKS) ciphertexts and ModSwitch Q -> q
        const size_t m = std::min<size_t>(cnt, 8);
        const NativeInteger qKS = lwe->GetqKS();
        std::uniform_int_distribution<uint64_t> uk(0, qKS.ConvertToInt() - 1);
        std::vector<LWECiphertext> in(m), gpu, cpu(m);
        for (size_t i = 0; i < m; ++i) {
            NativeVector a(N, qKS);
            for (uint32_t k = 0; k < N; ++k) a[k] = NativeInteger(uk(rng));
            in[i] = std::make_shared<LWECiphertextImpl>(std::move(a), NativeInteger(uk(rng)));
        }
        be->KeySwitchBatch(lwe, in, cc.GetSwitchKey(), gpu);
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.GetLWEScheme()->KeySwitch(lwe, cc.GetSwitchKey(), in[i]);
        report("Backend_KeySwitchBatch", m, count_diff(gpu, cpu));

        std::vector<LWECiphertext> inq(m), gq, cq(m);
        for (size_t i = 0; i < m; ++i) {
            NativeVector a(n, Q);
            for (uint32_t k = 0; k < n; ++k) a[k] = NativeInteger(uq(rng));
            inq[i] = std::make_shared<LWECiphertextImpl>(std::move(a), NativeInteger(uq(rng)));
        }
        be->ModSwitchBatch(lwe, inq, gq);
        for (size_t i = 0; i < m; ++i) cq[i] = cc.GetLWEScheme()->ModSwitch(lwe->Getq(), inq[i]);
        report("Backend_ModSwitchBatch", m, count_diff(gq, cq));
    }
    {
        // packed formats: key and ciphertext round trips through device memory
        auto buf = be->PackBootstrappingKey(cc.GetRefreshKey());
        RingGSWACCKey back;
        be->UnpackBootstrappingKey(buf, back);
        be->Free(buf);
        size_t bad = 0, seen = 0;
        const auto& e0 = cc.GetRefreshKey()->GetElements();
        const auto& e1 = back->GetElements();
        for (size_t i = 0; i < e0.size() && i < e1.size(); ++i)
            for (size_t j = 0; j < e0[i].size() && j < e1[i].size(); ++j)
                for (size_t k = 0; k < e0[i][j].size() && k < e1[i][j].size(); ++k) {
                    if (!e0[i][j][k] || !e1[i][j][k]) continue;
                    ++seen;
                    bad += e0[i][j][k]->GetElements() != e1[i][j][k]->GetElements();
                }
        report("Backend_PackUnpackBSK", seen, bad + (seen == 0));
        auto cb = be->PackCiphertexts(c1);
        std::vector<LWECiphertext> rt;
        be->UnpackCiphertexts(cb, rt);
        be->Free(cb);
        report("Backend_PackUnpackCiphertexts", cnt, count_diff(rt, c1));
    }
    // ---- reference edge cases at the seam (binfhe-base-scheme.cpp:81-93) -------------------------------------------
    {
        // (a) "extended" inputs — dimension N, modulus Q — are switched down first (SwitchCTtoqn), per operand
        const size_t m = std::min<size_t>(cnt, 6);
        std::vector<LWECiphertext> x(m), y(c2.begin(), c2.begin() + m), gpu, cpu(m);
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) x[i] = cc.EvalBinGate(OR, c1[i], c3[i], true);   // extended result of the reference
        auto r = EvalBinGateBatch(cc, NAND, x, y, gpu);
        size_t bad = r.success ? 0 : m;
        if (!r.success) std::printf("extended ct1: %s\n", r.error.c_str());
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.EvalBinGate(NAND, x[i], y[i]);
        if (r.success) bad += count_diff(gpu, cpu);
        r = EvalBinGateBatch(cc, XOR, y, x, gpu);                                      // extended ct2
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.EvalBinGate(XOR, y[i], x[i]);
        bad += r.success ? count_diff(gpu, cpu) : m;
        // mixed batch: extended and SMALL_DIM ct1 in one call
        std::vector<LWECiphertext> mix(m);
        for (size_t i = 0; i < m; ++i) mix[i] = (i & 1) ? x[i] : c1[i];
        r = EvalBinGateBatch(cc, AND, mix, y, gpu);
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.EvalBinGate(AND, mix[i], y[i]);
        bad += r.success ? count_diff(gpu, cpu) : m;
        report("EvalBinGateBatch_extended_inputs", 3 * m, bad);
        // (b) the same ciphertext object on both sides is refused with the reference's message
        std::vector<LWECiphertext> o;
        r = EvalBinGateBatch(cc, AND, c1, c1, o);
        report("EvalBinGateBatch_same_ct_refused", 1, (!r.success && r.error.find("independant") != std::string::npos) ? 0 : 1);
        // (c) a ct1 that claims plaintext modulus 8: the reference bootstraps a COPY of it, and the copy constructor drops m_p
        // (lwe-ciphertext.h:56), so the result is that of p = 4 — and Bootstrap() mixes both (table with 4, offset with p)
        std::vector<LWECiphertext> p8(m);
        for (size_t i = 0; i < m; ++i) { p8[i] = std::make_shared<LWECiphertextImpl>(*c1[i]); p8[i]->SetptModulus(NativeInteger(8)); }
        r = EvalBinGateBatch(cc, OR, p8, y, gpu);
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.EvalBinGate(OR, p8[i], y[i]);
        size_t badp = r.success ? count_diff(gpu, cpu) : m;
        r = BootstrapBatch(cc, p8, gpu);
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.Bootstrap(p8[i]);
        badp += r.success ? count_diff(gpu, cpu) : m;
        report("EvalBinGateBatch_ptmodulus", 2 * m, badp);
    }
    // ---- two async futures in flight on one backend (batch.cpp:320-349 runs them on std::async threads) -----------
    {
        std::vector<LWECiphertext> g1, g2, cpu1(cnt), cpu2(cnt);
        auto f1 = BootstrapBatchAsync(cc, c1, g1);
        auto f2 = BootstrapBatchAsync(cc, c2, g2);
        auto r1 = f1.Wait(), r2 = f2.Wait();
#pragma omp parallel for
        for (size_t i = 0; i < cnt; ++i) { cpu1[i] = cc.Bootstrap(c1[i]); cpu2[i] = cc.Bootstrap(c2[i]); }
        report("Async_overlapping_futures", 2 * cnt, (r1.success && r2.success) ? count_diff(g1, cpu1) + count_diff(g2, cpu2) : 2 * cnt);
    }
    // ---- e2e through the reference's own call: EvalBinGateBatch(cc, AND, vector<LWECiphertext> ...) -----------------
    if (e2e_gates) {
        std::vector<LWECiphertext> x(e2e_gates), y(e2e_gates), o;
        std::vector<int> mx(e2e_gates), my(e2e_gates);
#pragma omp parallel for
        for (size_t i = 0; i < e2e_gates; ++i) {
            mx[i] = (i * 7 + 1) & 1; my[i] = (i * 13 + 5) >> 2 & 1;
            x[i] = cc.Encrypt(sk, mx[i]); y[i] = cc.Encrypt(sk, my[i]);
        }
        EvalBinGateBatch(cc, AND, x, y, o);   // warm-up (staging buffers, kernels)
        double best = 1e30;
        backend::BackendB200::Timing tm;
        for (int rep = 0; rep < 3; ++rep) {
            o.clear();
            const double t0 = wall_ms();
            auto r = EvalBinGateBatch(cc, AND, x, y, o);
            const double dt = wall_ms() - t0;
            if (!r.success) { std::printf("e2e EvalBinGateBatch failed: %s\n", r.error.c_str()); return 2; }
            if (dt < best) { best = dt; tm = be->LastTiming(); }
        }
        size_t bad = 0;
        for (size_t i = 0; i < e2e_gates; ++i) { LWEPlaintext d; cc.Decrypt(sk, o[i], &d); bad += d != (mx[i] & my[i]); }
        std::vector<LWECiphertext> cpu(std::min<size_t>(e2e_gates, 64));
#pragma omp parallel for
        for (size_t i = 0; i < cpu.size(); ++i) cpu[i] = cc.EvalBinGate(AND, x[i], y[i]);
        for (size_t i = 0; i < cpu.size(); ++i) bad += !same_ct(o[i], cpu[i]);
        report("e2e_cpp_correct", e2e_gates, bad);
        std::printf("{\"e2e_cpp\": {\"gates\": %zu, \"devices\": %zu, \"ms\": %.2f, \"bootstraps_per_s\": %.1f, \"pack_ms\": %.2f, "
                    "\"call_ms\": %.2f, \"unpack_ms\": %.2f}}\n", e2e_gates, be->NumDevices(), best, e2e_gates / (best * 1e-3),
                    tm.pack_ms, tm.call_ms, tm.unpack_ms);
    }
    // ---- keys follow the context: a second BTKeyGen and a second BinFHEContext after RegisterB200 ------------------
    {
        auto sk2 = cc.KeyGen();
        cc.ClearBTKeys();                                  // (a second BTKeyGen alone re-uses the cached map entry)
        cc.BTKeyGen(sk2);                                  // the context now holds NEW key objects
        const size_t m = std::min<size_t>(cnt, 6);
        std::vector<LWECiphertext> x(m), y(m), gpu, cpu(m);
        for (size_t i = 0; i < m; ++i) { x[i] = cc.Encrypt(sk2, m1[i]); y[i] = cc.Encrypt(sk2, m2[i]); }
        auto r = EvalBinGateBatch(cc, NAND, x, y, gpu);
#pragma omp parallel for
        for (size_t i = 0; i < m; ++i) cpu[i] = cc.EvalBinGate(NAND, x[i], y[i]);
        size_t bad = r.success ? count_diff(gpu, cpu) : m;
        for (size_t i = 0; i < m && r.success; ++i) { LWEPlaintext d; cc.Decrypt(sk2, gpu[i], &d); bad += d != (1 - (m1[i] & m2[i])); }
        report("Rekey_follows_context", m, bad);

        BinFHEContext cc2;                                 // another parameter set in the same process
        cc2.GenerateBinFHEContext(TOY, ps == TOY && method == GINX ? AP : GINX);
        auto skb = cc2.KeyGen();
        cc2.BTKeyGen(skb);
        std::vector<LWECiphertext> u(m), v(m), g2, c2cpu(m);
        for (size_t i = 0; i < m; ++i) { u[i] = cc2.Encrypt(skb, m1[i]); v[i] = cc2.Encrypt(skb, m3[i]); }
        r = EvalBinGateBatch(cc2, OR, u, v, g2);
        for (size_t i = 0; i < m; ++i) c2cpu[i] = cc2.EvalBinGate(OR, u[i], v[i]);
        bad = r.success ? count_diff(g2, c2cpu) : m;
        r = EvalBinGateBatch(cc, NAND, x, y, gpu);         // and back to the first context
        bad += r.success ? count_diff(gpu, cpu) : m;
        report("Second_context_rebinds", 2 * m, bad);
    }
    if (dcrt) dcrt_check(65536, 32, 60, 2);
    std::printf("{\"host_parity\": \"%s\", \"devices\": %zu, \"failed_checks\": %d, %s}\n", set.c_str(), be->NumDevices(), g_fail,
                g_json.c_str());
    return g_fail ? 1 : 0;
}
