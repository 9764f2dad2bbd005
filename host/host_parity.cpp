// host_parity.cpp — drop-in check of the C++ host layer: the reference's own BinFHEContext generates parameters,
// keys and ciphertexts; the batch API (host/batch_b200.cpp) and the backend seam (host/backend_b200.cpp) run on the
// B200; every output is compared coefficient by coefficient with what the reference's CPU code returns for the same
// inputs (cc.EvalBinGate / Bootstrap / EvalFunc / EvalAcc / KeySwitch / ModSwitch).  TEST INFRASTRUCTURE: links the
// reference library built by oracle/Makefile.  usage: host_parity <TOY|STD128|STD128_AP|STD128_LMKCDEY> [count] [pairs]
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <random>
#include <string>

#include "backend_b200.h"
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

int main(int argc, char** argv) {
    const std::string set = argc > 1 ? argv[1] : "STD128";
    const size_t cnt = argc > 2 ? std::strtoul(argv[2], nullptr, 10) : 8;
    const size_t pairs = argc > 3 ? std::strtoul(argv[3], nullptr, 10) : 2;
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

    BinFHEContext cc;
    cc.GenerateBinFHEContext(ps, method);
    auto sk = cc.KeyGen();
    cc.BTKeyGen(sk);
    auto* be = backend::RegisterB200(cc, 0);
    std::printf("backend: %s, device memory %.1f GB, default=%s\n", be->Name().c_str(), be->DeviceMemory() / 1e9,
                backend::BackendTypeName(backend::CurrentBackend()->Type()).c_str());

    std::mt19937 rng(12345);
    std::vector<int> m1(cnt), m2(cnt), m3(cnt);
    std::vector<LWECiphertext> c1(cnt), c2(cnt), c3(cnt);
    for (size_t i = 0; i < cnt; ++i) {
        m1[i] = rng() & 1; m2[i] = rng() & 1; m3[i] = rng() & 1;
        c1[i] = cc.Encrypt(sk, m1[i]); c2[i] = cc.Encrypt(sk, m2[i]); c3[i] = cc.Encrypt(sk, m3[i]);
    }

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

    // ---- BatchDAG with an EvalFunc node: not a gate circuit, so Execute levels it on the host (one batched call per level
    // and operation kind) — same results as the reference's sequential walk
    if (!dag_lut.empty()) {
        auto build = [&](BatchDAG& d, std::vector<size_t>& outs, const std::vector<LWECiphertext>& fin) {
            for (size_t i = 0; i + 1 < cnt; i += 2) {
                size_t a = d.AddCiphertext(c1[i]), b = d.AddCiphertext(c2[i + 1]), e = d.AddCiphertext(fin[i]);
                size_t x = d.AddBinGate(NAND, a, b), f = d.AddEvalFunc(e, dag_lut), y = d.AddBinGate(XNOR, x, b);
                size_t u = d.AddBootstrap(y);
                outs.insert(outs.end(), {x, f, y, u});
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
    std::printf("{\"host_parity\": \"%s\", \"failed_checks\": %d, %s}\n", set.c_str(), g_fail, g_json.c_str());
    return g_fail ? 1 : 0;
}
