// packed_parity.cpp — TEST binary for the packed key wire formats (host/packed_b200.cpp).
// CPU part (always): reference keygen -> PackBootstrappingKey / PackSwitchingKey -> Unpack* -> every coefficient equal,
// malformed input rejected, and the reference's own CPU EvalBinGate on a context loaded with the UNPACKED keys gives
// the same ciphertext as on the original keys.  GPU part (argv[3] = 1): a fresh context with no keys of its own, bound
// to the B200 and fed BackendB200::LoadPacked(bytes), evaluates a gate batch identical to the reference's CPU results.
// usage: packed_parity <TOY|TOY_AP|TOY_LMKCDEY|STD128|...> <count> <gpu 0|1>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>

#include "backend_b200.h"
#include "batch/binfhe-batch.h"
#include "packed_b200.h"

using namespace lux::fhe;

static bool same_ct(const LWECiphertext& x, const LWECiphertext& y) {
    return x->GetA() == y->GetA() && x->GetB() == y->GetB() && x->GetModulus() == y->GetModulus();
}

int main(int argc, char** argv) {
    const std::string set = argc > 1 ? argv[1] : "TOY";
    const size_t count = argc > 2 ? std::stoul(argv[2]) : 8;
    const bool gpu = argc > 3 && std::stoi(argv[3]) != 0;
    const std::map<std::string, std::pair<BINFHE_PARAMSET, BINFHE_METHOD>> sets = {
        {"TOY", {TOY, GINX}}, {"TOY_AP", {TOY, AP}}, {"TOY_LMKCDEY", {TOY, LMKCDEY}},
        {"STD128", {STD128, GINX}}, {"STD128_AP", {STD128_AP, AP}}, {"STD128_LMKCDEY", {STD128_LMKCDEY, LMKCDEY}}};
    auto it = sets.find(set);
    if (it == sets.end()) { std::fprintf(stderr, "unknown set %s\n", set.c_str()); return 2; }
    int fails = 0;
    auto expect = [&](bool ok, const char* what) { std::printf("%-44s %s\n", what, ok ? "ok" : "FAILED"); fails += !ok; };

    BinFHEContext cc;
    cc.GenerateBinFHEContext(it->second.first, it->second.second);
    auto sk = cc.KeyGen();
    cc.BTKeyGen(sk);
    auto rg = cc.GetParams()->GetRingGSWParams();
    auto lwe = cc.GetParams()->GetLWEParams();

    const std::vector<uint8_t> pb = backend::PackBootstrappingKey(cc.GetRefreshKey(), rg);
    const std::vector<uint8_t> pk = backend::PackSwitchingKey(cc.GetSwitchKey(), lwe);
    std::printf("packed BSK %.1f MB, KSK %.1f MB\n", pb.size() / 1e6, pk.size() / 1e6);
    expect(backend::ValidatePackedHeader(pb.data(), pb.size()) && backend::GetPackedType(pb.data(), pb.size()) == backend::PackedType::BOOTSTRAPPING_KEY, "BSK header");
    expect(backend::ValidatePackedHeader(pk.data(), pk.size()) && backend::GetPackedType(pk.data(), pk.size()) == backend::PackedType::SWITCHING_KEY, "KSK header");

    RingGSWACCKey ek2 = backend::UnpackBootstrappingKey(pb.data(), pb.size(), rg);
    LWESwitchingKey ks2 = backend::UnpackSwitchingKey(pk.data(), pk.size(), lwe);
    {   // every slot and coefficient
        const auto& a = cc.GetRefreshKey()->GetElements();
        const auto& b = ek2->GetElements();
        bool ok = a.size() == b.size();
        size_t present = 0;
        for (size_t i = 0; ok && i < a.size(); ++i) {
            ok = a[i].size() == b[i].size();
            for (size_t j = 0; ok && j < a[i].size(); ++j) {
                ok = a[i][j].size() == b[i][j].size();
                for (size_t k = 0; ok && k < a[i][j].size(); ++k) {
                    if (!a[i][j][k] || !b[i][j][k]) { ok = !a[i][j][k] && !b[i][j][k]; continue; }
                    ++present;
                    ok = a[i][j][k]->GetElements() == b[i][j][k]->GetElements();
                }
            }
        }
        std::printf("  %zu RGSW entries compared\n", present);
        expect(ok && present > 0, "BSK pack -> unpack identical");
        expect(*cc.GetSwitchKey() == *ks2, "KSK pack -> unpack identical");
    }
    {   // malformed inputs are refused, not read
        bool threw = false;
        try { backend::UnpackBootstrappingKey(pb.data(), pb.size() - 8, rg); } catch (const std::exception&) { threw = true; }
        expect(threw, "truncated BSK rejected");
        threw = false;
        try { backend::UnpackBootstrappingKey(pk.data(), pk.size(), rg); } catch (const std::exception&) { threw = true; }
        expect(threw, "KSK bytes rejected as BSK");
        std::vector<uint8_t> bad = pk;
        bad[0] ^= 0xFF;
        threw = false;
        try { backend::UnpackSwitchingKey(bad.data(), bad.size(), lwe); } catch (const std::exception&) { threw = true; }
        expect(threw, "bad magic rejected");
        expect(backend::PackBootstrappingKey(nullptr, rg).empty() && backend::PackSwitchingKey(nullptr, lwe).empty(), "empty key -> empty bytes");
    }
    // LWE formats come from the reference's own packed.cpp compiled into this library
    std::vector<LWECiphertext> a(count), b(count), ref(count);
    for (size_t i = 0; i < count; ++i) { a[i] = cc.Encrypt(sk, i & 1); b[i] = cc.Encrypt(sk, (i >> 1) & 1); }
    {
        auto bytes = backend::PackLWEBatch(a);
        auto back = backend::UnpackLWEBatch(bytes.data(), bytes.size());
        bool ok = back.size() == count;
        // the reference's format carries values only (log_q = 64, q = 0: packed.cpp:98-99), so compare values, not moduli
        for (size_t i = 0; ok && i < count; ++i) {
            ok = back[i]->GetB() == a[i]->GetB() && back[i]->GetA().GetLength() == a[i]->GetA().GetLength();
            for (uint32_t t = 0; ok && t < a[i]->GetA().GetLength(); ++t) ok = back[i]->GetA()[t] == a[i]->GetA()[t];
        }
        expect(ok, "PackLWEBatch -> UnpackLWEBatch (values)");
    }
    for (size_t i = 0; i < count; ++i) ref[i] = cc.EvalBinGate(NAND, a[i], b[i]);

    // a second context that never ran keygen: parameters + unpacked keys only
    BinFHEContext cc2;
    cc2.GenerateBinFHEContext(it->second.first, it->second.second);
    cc2.BTKeyLoad({ek2, ks2});
    {
        bool ok = true;
        for (size_t i = 0; ok && i < std::min<size_t>(count, 2); ++i) ok = same_ct(cc2.EvalBinGate(NAND, a[i], b[i]), ref[i]);
        expect(ok, "CPU gate on unpacked keys == original keys");
    }
    if (gpu) {
        BinFHEContext cc3;   // no keys at all on the host side
        cc3.GenerateBinFHEContext(it->second.first, it->second.second);
        try {
            auto* be = backend::RegisterB200(cc3, 0);
            be->LoadPacked(pb.data(), pb.size(), pk.data(), pk.size());
            std::vector<LWECiphertext> out;
            const BatchResult r = EvalBinGateBatch(cc3, NAND, a, b, out);
            bool ok = r.success && out.size() == count;
            for (size_t i = 0; ok && i < count; ++i) ok = same_ct(out[i], ref[i]);
            expect(ok, "B200 gates on LoadPacked keys == reference");
            bool threw = false;
            try { be->LoadPacked(pb.data(), pb.size() - 8, nullptr, 0); } catch (const std::exception&) { threw = true; }
            expect(threw, "LoadPacked rejects truncated bytes");
        } catch (const std::exception& e) {
            std::printf("B200: %s\n", e.what());
            ++fails;
        }
    }
    std::printf("{\"packed_parity\": \"%s\", \"failed_checks\": %d}\n", set.c_str(), fails);
    return fails ? 1 : 0;
}
