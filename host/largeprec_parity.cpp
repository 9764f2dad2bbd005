// largeprec_parity.cpp — TEST driver: the reference's large-precision context (binfhecontext.cpp:53-104) generates keys
// (one key set, or three with timeOptimization) and ciphertexts at Q = 2^logQ; host/largeprec_b200 runs EvalFloorBatch /
// EvalSignBatch / EvalDecompBatch on the B200; every output ciphertext is compared with cc.EvalFloor / EvalSign / EvalDecomp
// on the CPU (the cases of src/binfhe/unittest/UnitTestFunc.cpp:45-266: messages around p/2) and decrypted.
// usage: largeprec_parity <TOY|STD128> <logQ> <timeOpt 0|1> [count]
#include <cstdio>
#include <cstdlib>
#include <string>

#include "backend_b200.h"
#include "largeprec_b200.h"

using namespace lux::fhe;

static int g_fail = 0;
static std::string g_json;
static void report(const char* name, size_t n, size_t bad) {
    std::printf("%-24s %4zu compared, %zu mismatches\n", name, n, bad);
    if (bad) ++g_fail;
    g_json += std::string(g_json.empty() ? "" : ", ") + "\"" + name + "\": " + (bad ? "false" : "true");
}
static bool same_ct(const LWECiphertext& x, const LWECiphertext& y) {
    return x && y && x->GetLength() == y->GetLength() && x->GetModulus() == y->GetModulus() && x->GetB() == y->GetB() &&
           x->GetA() == y->GetA();
}

int main(int argc, char** argv) {
    const std::string set = argc > 1 ? argv[1] : "TOY";
    const uint32_t logQ = argc > 2 ? std::strtoul(argv[2], nullptr, 10) : 17;
    const bool timeOpt = argc > 3 ? std::atoi(argv[3]) != 0 : true;
    const size_t cnt = argc > 4 ? std::strtoul(argv[4], nullptr, 10) : 8;
    BinFHEContext cc;
    cc.GenerateBinFHEContext(set == "STD128" ? STD128 : TOY, false, logQ, 0, GINX, timeOpt);
    auto sk = cc.KeyGen();
    cc.BTKeyGen(sk);
    if (luxb200_device_count() < 1) { std::printf("no CUDA device: there is no CPU fallback\n"); return 3; }
    const uint64_t Q = uint64_t(1) << logQ, q = cc.GetParams()->GetLWEParams()->Getq().ConvertToInt<uint64_t>();
    const uint64_t p = cc.GetMaxPlaintextSpace().ConvertToInt<uint64_t>() * (Q / q);
    std::printf("large-precision context: n=%u N=%u q=%llu Q=2^%u p=%llu, %zu key set(s)\n", cc.GetParams()->GetLWEParams()->Getn(),
                cc.GetParams()->GetLWEParams()->GetN(), (unsigned long long)q, logQ, (unsigned long long)p, cc.GetBTKeyMap()->size());
    std::vector<LWECiphertext> in(cnt);
    std::vector<int64_t> msg(cnt);
    for (size_t i = 0; i < cnt; ++i) {
        msg[i] = int64_t(p / 2) + int64_t(i) - int64_t(cnt / 2);
        in[i]  = cc.Encrypt(sk, msg[i], FRESH, p, Q);
    }
    {
        std::vector<LWECiphertext> gpu, cpu(cnt);
        auto r = EvalFloorBatch(cc, in, gpu);
        if (!r.success) { std::printf("EvalFloorBatch failed: %s\n", r.error.c_str()); return 2; }
        for (size_t i = 0; i < cnt; ++i) cpu[i] = cc.EvalFloor(in[i]);
        size_t bad = 0;
        for (size_t i = 0; i < cnt; ++i) bad += !same_ct(gpu[i], cpu[i]);
        report("EvalFloorBatch", cnt, bad);
    }
    for (int ss = 0; ss < 2; ++ss) {
        std::vector<LWECiphertext> gpu, cpu(cnt);
        auto r = EvalSignBatch(cc, in, gpu, ss != 0);
        if (!r.success) { std::printf("EvalSignBatch failed: %s\n", r.error.c_str()); return 2; }
        for (size_t i = 0; i < cnt; ++i) cpu[i] = cc.EvalSign(in[i], ss != 0);
        size_t bad = 0;
        for (size_t i = 0; i < cnt; ++i) bad += !same_ct(gpu[i], cpu[i]);
        if (ss == 0)
            for (size_t i = 0; i < cnt; ++i) {
                LWEPlaintext d;
                cc.Decrypt(sk, gpu[i], &d, 2);
                bad += d != (msg[i] >= int64_t(p / 2) ? 1 : 0);
            }
        report(ss ? "EvalSignBatch_schemeSwitch" : "EvalSignBatch", cnt, bad);
    }
    {
        std::vector<std::vector<LWECiphertext>> gpu;
        auto r = EvalDecompBatch(cc, in, gpu);
        if (!r.success) { std::printf("EvalDecompBatch failed: %s\n", r.error.c_str()); return 2; }
        size_t bad = 0, seen = 0;
        for (size_t i = 0; i < cnt; ++i) {
            auto cpu = cc.EvalDecomp(in[i]);
            bad += cpu.size() != gpu[i].size();
            for (size_t d = 0; d < cpu.size() && d < gpu[i].size(); ++d, ++seen) bad += !same_ct(gpu[i][d], cpu[d]);
        }
        report("EvalDecompBatch", seen, bad);
    }
    {
        // small-precision input is refused with the reference's message
        std::vector<LWECiphertext> small{cc.Encrypt(sk, 1)}, out;
        auto r = EvalSignBatch(cc, small, out);
        report("EvalSignBatch_refuses_small", 1, (!r.success && r.error.find("only for large precision") != std::string::npos) ? 0 : 1);
    }
    ReleaseLargePrecisionB200();
    std::printf("{\"largeprec_parity\": \"%s\", \"logQ\": %u, \"key_sets\": %zu, \"failed_checks\": %d, %s}\n", set.c_str(), logQ,
                cc.GetBTKeyMap()->size(), g_fail, g_json.c_str());
    return g_fail ? 1 : 0;
}
