// api.cu — C ABI of libluxb200.so (see include/luxb200.h for the reference interfaces each entry replaces).
#include <algorithm>

#include "api_common.h"
#include "aux_kernels.cuh"
#include "ks_tile.cuh"
#include "br_sched.cuh"
#include "br_wide.cuh"
#include "keygen.cuh"
#include "ntt_kernels.cuh"

thread_local std::string g_last_error;

int lb_fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}
int lb_set_device(int device) {
    int cnt = 0;
    cudaError_t e = cudaGetDeviceCount(&cnt);
    if (e != cudaSuccess || cnt == 0)
        return lb_fail(LUXB200_ERR_NO_DEVICE, "no CUDA device available (%s); libluxb200 has no CPU fallback",
                       e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= cnt) return lb_fail(LUXB200_ERR_INVALID_ARG, "device %d out of range [0,%d)", device, cnt);
    LB_CUDA(cudaSetDevice(device));
    return 0;
}
static inline unsigned nblk(uint64_t total, unsigned bs) { return (unsigned)((total + bs - 1) / bs); }

// ---- u32 NTT tables -------------------------------------------------------------------------------------------
int Ntt32Tables::build(uint32_t N_, uint64_t Q_, uint64_t psi_) {
    using namespace lb;
    std::vector<uint64_t> tw, twi;
    ntt_tables(N_, Q_, psi_, tw, twi);
    psi = psi_;
    return build_from(N_, Q_, tw, twi, invmod_prime(N_ % Q_, Q_));
}
// Half h of a 2N-point transform as an N-point one (br_quad.cuh): after the first Cooley-Tukey stage the two halves are
// independent N-point transforms whose stage-m twiddles are the 2N-point table's entries 2m + h m + i; ninv2 = (2N)^-1
int Ntt32Tables::build_half(uint32_t N_, uint64_t Q_, const std::vector<uint64_t>& tw2, const std::vector<uint64_t>& twi2,
                            uint32_t h, uint64_t ninv2) {
    std::vector<uint64_t> tw(N_, 1), twi(N_, 1);
    for (uint32_t m = 1; m < N_; m <<= 1)
        for (uint32_t i = 0; i < m; ++i) {
            tw[m + i]  = tw2[2 * m + h * m + i];
            twi[m + i] = twi2[2 * m + h * m + i];
        }
    return build_from(N_, Q_, tw, twi, ninv2);
}
int Ntt32Tables::build_from(uint32_t N_, uint64_t Q_, const std::vector<uint64_t>& tw, const std::vector<uint64_t>& twi, uint64_t ninv) {
    using namespace lb;
    N = N_; Q = Q_;
    logN = ilog2(N);
    if (Q >= (1ull << 30)) return lb_fail(LUXB200_ERR_UNSUPPORTED, "u32 NTT path needs Q < 2^30 (got %llu)", (unsigned long long)Q);
    logH = 32 - bits_of(Q);
    if (logH > 5) logH = 5;
    const uint32_t E = N / 32, TB = (E / 2 < 16) ? E / 2 : 16;
    memset(&C, 0, sizeof(C));
    for (uint32_t i = 1; i < E && i < 64; ++i) {
        C.twA[i]  = make_uint2((u32)tw[i], shoup32(tw[i], Q));
        C.twiA[i] = make_uint2((u32)twi[i], shoup32(twi[i], Q));
    }
    C.Q = (u32)Q; C.Q2 = (u32)(2 * Q); C.mu32 = (u32)((1ull << 32) / Q);
    for (int k = 0; k < 8; ++k) C.QM[k] = (u32)(Q << k);   // only multiples below 2^32 are ever used (bounds <= 2^LOGH)
    C.ninv = (u32)ninv; C.ninv_s = shoup32(ninv, Q);
    uint64_t w1 = mulmod(twi[1], ninv, Q);
    C.w1ninv = (u32)w1; C.w1ninv_s = shoup32(w1, Q);
    // lane-major phase-B tables
    std::vector<uint2> hB, hBi, hF(N), hFi(N);
    uint32_t off = 0;
    for (uint32_t t = TB; t >= 1; t >>= 1) {
        const uint32_t g = E / (2 * t);
        hB.resize((off + g) * 32);
        hBi.resize((off + g) * 32);
        for (uint32_t i = 0; i < g; ++i)
            for (uint32_t lane = 0; lane < 32; ++lane) {
                const uint32_t idx = N / (2 * t) + g * lane + i;
                hB[(off + i) * 32 + lane]  = make_uint2((u32)tw[idx], shoup32(tw[idx], Q));
                hBi[(off + i) * 32 + lane] = make_uint2((u32)twi[idx], shoup32(twi[idx], Q));
            }
        off += g;
    }
    for (uint32_t i = 0; i < N; ++i) {
        hF[i]  = make_uint2((u32)tw[i], shoup32(tw[i], Q));
        hFi[i] = make_uint2((u32)twi[i], shoup32(twi[i], Q));
    }
    LB_TRY(twB.ensure(hB.size() * sizeof(uint2)));
    LB_TRY(twiB.ensure(hBi.size() * sizeof(uint2)));
    LB_TRY(twfull.ensure(N * sizeof(uint2)));
    LB_TRY(twifull.ensure(N * sizeof(uint2)));
    LB_CUDA(cudaMemcpy(twB.p, hB.data(), hB.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    LB_CUDA(cudaMemcpy(twiB.p, hBi.data(), hBi.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    LB_CUDA(cudaMemcpy(twfull.p, hF.data(), N * sizeof(uint2), cudaMemcpyHostToDevice));
    LB_CUDA(cudaMemcpy(twifull.p, hFi.data(), N * sizeof(uint2), cudaMemcpyHostToDevice));
    return 0;
}

// (w, Shoup) twiddle tables, psi powers and reduction constants of the shared-memory kernels (br_wide.cuh, keygen.cuh);
// W.wsize must be set (4: Q < 2^30, 8 otherwise)
int lb_build_wide_tables(luxb200_ctx* c) {
    using namespace lb;
    luxb200_ctx::Wide& W = c->wide;
    if (W.tw.p) return 0;
    const luxb200_params* p = &c->p;
    const uint32_t N = p->N;
        const uint64_t Q = p->Q;
        std::vector<uint64_t> tw, twi, pw(2 * (size_t)N);
        ntt_tables(N, Q, c->p.psi, tw, twi);
        pw[0] = 1;
        for (uint32_t i = 1; i < 2 * N; ++i) pw[i] = mulmod(pw[i - 1], c->p.psi, Q);
        W.ninv = invmod_prime(N % Q, Q);
    int rcw = 0;
        if (W.wsize == 8) {
            W.ninv_s = shoup64(W.ninv, Q);
            W.mu = (uint64_t)((((u128_t)1) << 64) / Q);
            W.r = (uint64_t)((((u128_t)1) << 64) % Q);
            W.rs = shoup64(W.r, Q);
            std::vector<ulonglong2> t(N), ti(N);
            for (uint32_t i = 0; i < N; ++i) { t[i] = make_ulonglong2(tw[i], shoup64(tw[i], Q)); ti[i] = make_ulonglong2(twi[i], shoup64(twi[i], Q)); }
            rcw = W.tw.ensure(N * 16); if (!rcw) rcw = W.twi.ensure(N * 16); if (!rcw) rcw = W.pw.ensure(2 * (size_t)N * 8);
            if (!rcw && (cudaMemcpy(W.tw.p, t.data(), N * 16, cudaMemcpyHostToDevice) != cudaSuccess ||
                         cudaMemcpy(W.twi.p, ti.data(), N * 16, cudaMemcpyHostToDevice) != cudaSuccess ||
                         cudaMemcpy(W.pw.p, pw.data(), 2 * (size_t)N * 8, cudaMemcpyHostToDevice) != cudaSuccess))
                rcw = lb_fail(LUXB200_ERR_CUDA, "table upload failed");
        } else {
            W.ninv_s = shoup32(W.ninv, Q);
            W.mu = (uint64_t)((1ull << 32) / Q);
            W.r = (uint64_t)((1ull << 32) % Q);
            W.rs = shoup32(W.r, Q);
            std::vector<uint2> t(N), ti(N);
            std::vector<u32> pw32(2 * (size_t)N);
            for (uint32_t i = 0; i < N; ++i) { t[i] = make_uint2((u32)tw[i], shoup32(tw[i], Q)); ti[i] = make_uint2((u32)twi[i], shoup32(twi[i], Q)); }
            for (uint32_t i = 0; i < 2 * N; ++i) pw32[i] = (u32)pw[i];
            rcw = W.tw.ensure(N * 8); if (!rcw) rcw = W.twi.ensure(N * 8); if (!rcw) rcw = W.pw.ensure(2 * (size_t)N * 4);
            if (!rcw && (cudaMemcpy(W.tw.p, t.data(), N * 8, cudaMemcpyHostToDevice) != cudaSuccess ||
                         cudaMemcpy(W.twi.p, ti.data(), N * 8, cudaMemcpyHostToDevice) != cudaSuccess ||
                         cudaMemcpy(W.pw.p, pw32.data(), 2 * (size_t)N * 4, cudaMemcpyHostToDevice) != cudaSuccess))
                rcw = lb_fail(LUXB200_ERR_CUDA, "table upload failed");
        }
    if (!rcw && cudaDeviceSynchronize() != cudaSuccess) rcw = lb_fail(LUXB200_ERR_CUDA, "table upload failed");   // the pageable uploads have landed
    return rcw;
}

extern "C" {

const char* luxb200_last_error(void) { return g_last_error.c_str(); }
int luxb200_version(void) { return 100; }
int luxb200_device_count(void) {
    int cnt = 0;
    if (cudaGetDeviceCount(&cnt) != cudaSuccess) return 0;
    return cnt;
}
uint64_t luxb200_last_prime(uint32_t bits, uint64_t m) { return lb::last_prime(bits, m); }
uint64_t luxb200_previous_prime(uint64_t q, uint64_t m) { return lb::previous_prime(q, m); }
uint64_t luxb200_root_of_unity(uint64_t m, uint64_t q) { return lb::root_of_unity(m, q); }

int luxb200_ctx_create(const luxb200_params* p, int device, luxb200_ctx** out) {
    using namespace lb;
    if (!p || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    if (p->N < 256 || (p->N & (p->N - 1))) return lb_fail(LUXB200_ERR_INVALID_ARG, "N must be a power of two >= 256");
    if (p->method < LUXB200_AP || p->method > LUXB200_LMKCDEY) return lb_fail(LUXB200_ERR_INVALID_ARG, "bad method %u", p->method);
    if (p->baseG == 0 || (p->baseG & (p->baseG - 1))) return lb_fail(LUXB200_ERR_INVALID_ARG, "Gadget base should be a power of two.");
    if (p->q == 0 || (p->q & (p->q - 1)) || (2ull * p->N) % p->q) return lb_fail(LUXB200_ERR_INVALID_ARG, "q must be a power of two dividing 2N");
    if (p->Q >= (1ull << 60) || !is_prime(p->Q) || (p->Q - 1) % (2ull * p->N))
        return lb_fail(LUXB200_ERR_UNSUPPORTED, "Q must be an NTT prime below 2^60");
    if (p->N > 4096) return lb_fail(LUXB200_ERR_UNSUPPORTED, "blind rotation kernels are built for N <= 4096 (got %u)", p->N);
    if (p->baseKS < 2) return lb_fail(LUXB200_ERR_INVALID_ARG, "baseKS must be at least 2");
    LB_TRY(lb_set_device(device));
    luxb200_ctx* c = new luxb200_ctx();
    c->p = *p;
    c->device = device;
    c->logN = ilog2(p->N);
    c->digitsG = digit_count(p->Q, p->baseG);
    c->digitsKS = digit_count(p->qKS, p->baseKS);
    c->digitsR = digit_count(p->q, p->baseR ? p->baseR : 2);
    c->rows = 2 * (c->digitsG - 1);
    if (c->p.psi == 0) c->p.psi = root_of_unity(2ull * p->N, p->Q);
    if (c->digitsG < 2 || c->digitsG > 9) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_UNSUPPORTED, "digitsG = %u not supported", c->digitsG); }
    // column sums of the key switch: N * digitsKS entries below qKS must fit the 64-bit accumulator
    if ((u128_t)p->qKS * p->N * c->digitsKS >= ((u128_t)1 << 64)) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_UNSUPPORTED, "qKS too large for the key-switch accumulator"); }
    // the register-resident warp-pair kernels cover N = 1024 with 2 or 3 digits and N = 512 with 2, Q < 2^30; everything else
    // (N = 2048, 64-bit Q, 1 / 4 / 5 digits) runs the shared-memory kernel of br_wide.cuh
    // r2: N = 2048 with 2 or 3 digits (GINX and AP) runs the same register-resident code as four warps per ciphertext
    // (br_quad.cuh)
    c->quad = p->Q < (1ull << 30) && p->Q >= (1ull << 24) && c->logN == 11 && (c->rows == 4 || c->rows == 6 || (c->rows == 8 && p->Q < (1ull << 27)));
    if (c->quad && bits_of(p->Q) < 30) {
        // below 30 bits these kernels use reduce64x, which leaves less than (xmax / 2^(s+32) + 2^s / Q + 1) Q: that must stay
        // under 4Q (it is 3.0 / 3.75 for the 29-bit sets with two / three digits); a modulus that leaves no margin runs wide
        const u128_t Q = p->Q;
        u128_t xmax = (u128_t)c->rows * 2 * Q * Q;
        if (8 * Q * Q + 4 * Q > xmax) xmax = 8 * Q * Q + 4 * Q;
        uint32_t b = 0;
        while ((xmax >> b) != 0) ++b;
        const int s = b > 32 ? (int)b - 32 : 0;
        const long double bound = (long double)xmax / std::ldexp(1.0L, s + 32) + std::ldexp(1.0L, s) / (long double)p->Q + 1.0L;
        if (bound > 3.999L) c->quad = false;
    }
    const bool fast = p->Q < (1ull << 30) && ((c->logN == 10 && (c->rows == 4 || c->rows == 6)) || (c->logN == 9 && c->rows == 4) || c->quad);
    c->wide.on = !fast;
    if (c->wide.on) {
        luxb200_ctx::Wide& W = c->wide;
        W.wsize = p->Q < (1ull << 30) ? 4 : 8;
        const uint32_t N = p->N;
        W.smem = (size_t)(2 + c->rows) * (N + N / 16) * W.wsize;
        if (W.smem > 232448 || N < 256) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_UNSUPPORTED, "N = %u with %u digit polynomials does not fit in shared memory", N, c->rows); }
        int rcw = lb_build_wide_tables(c);
        if (rcw) { luxb200_ctx_destroy(c); return rcw; }
    }
    int rc = 0;
    if (!c->wide.on) {
    // reduce64 constants: x < max(rows * 2Q * Q, 8 Q^2 + 4Q)
    {
        u128_t Q = p->Q, xmax = (u128_t)c->rows * 2 * Q * Q;
        u128_t alt = 8 * Q * Q + 4 * Q;
        if (alt > xmax) xmax = alt;
        uint32_t b = 0;
        while ((xmax >> b) != 0) ++b;
        uint32_t s = b > 32 ? b - 32 : 0;
        // mu = floor(2^(s+32) / Q) fits 32 bits and the truncated Barrett step keeps its error bound when 2^s < Q; the N = 2048
        // kernels reduce in two words instead (reduce64w: any x < 2^64) and carry R = 2^32 mod Q with its Shoup companion
        const u128_t mu = (((u128_t)1) << (s + 32)) / Q;
        if (c->quad) {
            if (b > 64) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_UNSUPPORTED, "Q too large for the 64-bit accumulation (rows=%u)", c->rows); }
            const uint64_t R = (uint64_t)((((u128_t)1) << 32) % Q);
            c->red_r = (uint32_t)R; c->red_rs = shoup32(R, p->Q);
            c->red_s = s; c->red_mu = (uint32_t)mu; c->red_mu_hi = (uint32_t)(mu >> 32);
        } else {
            if ((mu >> 32) != 0) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_UNSUPPORTED, "Q too large for the 64->32 reduction (rows=%u)", c->rows); }
            c->red_s = s;
            c->red_mu = (uint32_t)mu;
            c->red_mu_hi = 0;
        }
    }
    if (c->quad) {
        std::vector<uint64_t> tw2, twi2;
        ntt_tables(p->N, p->Q, c->p.psi, tw2, twi2);
        const uint64_t ninv2 = invmod_prime(p->N % p->Q, p->Q);
        rc = c->ntt.build_half(p->N / 2, p->Q, tw2, twi2, 0, ninv2);
        if (!rc) rc = c->ntt1.build_half(p->N / 2, p->Q, tw2, twi2, 1, ninv2);
        c->xw = (uint32_t)tw2[1]; c->xws = shoup32(tw2[1], p->Q);
        c->xwi = (uint32_t)twi2[1]; c->xwis = shoup32(twi2[1], p->Q);
    } else {
        rc = c->ntt.build(p->N, p->Q, c->p.psi);
    }
    }
    if (rc) { luxb200_ctx_destroy(c); return rc; }
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    for (int i = 0; i < 5 && e == cudaSuccess; ++i) e = cudaEventCreate(&c->ev[i]);
    if (e != cudaSuccess) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_CUDA, "stream/event creation failed: %s", cudaGetErrorString(e)); }
    if (p->method == LUXB200_GINX && !c->wide.on) {
        // monomials X^a - 1 in EVALUATION format (rgsw-cryptoparameters.cpp:96-113), sigma order:
        // value at index p is psi^(a (2 brv(p) + 1)) - 1
        const uint32_t N = p->N, M = 2 * N;
        const uint32_t E = (c->quad ? N / 2 : N) / 32, HM = c->quad ? N / 2 - 1 : N - 1;   // quad: [h][sigma order of p mod 1024]
        std::vector<uint64_t> pw(M);
        pw[0] = 1;
        for (uint32_t i = 1; i < M; ++i) pw[i] = mulmod(pw[i - 1], c->p.psi, p->Q);
        std::vector<u32> h((size_t)M * N);
        std::vector<uint32_t> odd(N);
        for (uint32_t pp = 0; pp < N; ++pp) odd[pp] = 2 * bitrev(pp, c->logN) + 1;
        for (uint32_t a = 0; a < M; ++a)
            for (uint32_t pp = 0; pp < N; ++pp) {
                const uint32_t ps = pp & HM, l = ps / E, r = ps % E;
                const uint64_t v = pw[((uint64_t)a * odd[pp]) % M];
                h[(size_t)a * N + (pp - ps) + (r >> 2) * 128 + 4 * l + (r & 3)] = (u32)(v ? v - 1 : p->Q - 1);
            }
        rc = c->mono.ensure(h.size() * 4);
        if (!rc && cudaMemcpy(c->mono.p, h.data(), h.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = lb_fail(LUXB200_ERR_CUDA, "mono upload failed");
        if (rc) { luxb200_ctx_destroy(c); return rc; }
    }
    if (p->method == LUXB200_LMKCDEY) {
        // EVALUATION-format automorphism maps (PrecomputeAutoMap, nbtheory2.cpp:264-275) for k = 2N-5, 5^1..5^w,
        // as gather indices in sigma order: out[sigma(p)] = in[sigma(map[p])]
        const uint32_t N = p->N, M = 2 * N, w = p->numAutoKeys;
        const uint32_t E = (c->quad ? N / 2 : N) / 32, HM = c->quad ? N / 2 - 1 : N - 1;   // quad: two halves, each in its own sigma order
        std::vector<u32> h((size_t)(w + 1) * N);
        for (uint32_t t = 0; t <= w; ++t) {
            uint32_t k = (t == 0) ? M - 5 : (uint32_t)powmod(5, t, M);
            for (uint32_t j = 0; j < N; ++j) {
                uint32_t jt = 2 * j + 1;
                uint32_t idx = (uint32_t)(((uint64_t)jt * k) % M) >> 1;
                uint32_t dst = bitrev(j, c->logN), src = bitrev(idx, c->logN);
                uint32_t ds = dst & HM, ld = ds / E, rd = ds % E;
                h[(size_t)t * N + (dst - ds) + (rd >> 2) * 128 + 4 * ld + (rd & 3)] = src;  // natural EVAL index of the source
            }
        }
        if (c->wide.on) {   // natural order for the shared-memory kernel: out[brv(j)] = in[brv(idx)]
            for (uint32_t t = 0; t <= w; ++t) {
                uint32_t k = (t == 0) ? M - 5 : (uint32_t)powmod(5, t, M);
                for (uint32_t j = 0; j < N; ++j) {
                    uint32_t idx = (uint32_t)(((uint64_t)(2 * j + 1) * k) % M) >> 1;
                    h[(size_t)t * N + bitrev(j, c->logN)] = bitrev(idx, c->logN);
                }
            }
        }
        DevBuf& amap = c->wide.on ? c->wide.automap : c->automap;
        rc = amap.ensure(h.size() * 4);
        if (!rc && cudaMemcpy(amap.p, h.data(), h.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = lb_fail(LUXB200_ERR_CUDA, "automap upload failed");
        if (rc) { luxb200_ctx_destroy(c); return rc; }
        // logGen (rgsw-cryptoparameters.cpp:115-127)
        std::vector<i32> lg(M, 0);
        uint32_t gp = 1;
        lg[M - gp] = (i32)M;
        for (uint32_t i = 1; i < N / 2; ++i) {
            gp = (gp * 5) % M;
            lg[gp] = (i32)i;
            lg[M - gp] = -(i32)i;
        }
        rc = c->loggen.ensure(M * 4);
        if (!rc && cudaMemcpy(c->loggen.p, lg.data(), M * 4, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = lb_fail(LUXB200_ERR_CUDA, "logGen upload failed");
        if (rc) { luxb200_ctx_destroy(c); return rc; }
    }
    // KSK element type
    c->ks_elem = (p->qKS <= 65536) ? 2 : (p->qKS <= (1ull << 32)) ? 4 : 8;
    c->npad = (p->n + 7) & ~7u;
    // the table uploads above were synchronous copies from pageable memory, which may return before their DMA has landed; the
    // context's non-blocking stream does not order itself behind them
    if (cudaDeviceSynchronize() != cudaSuccess) { luxb200_ctx_destroy(c); return lb_fail(LUXB200_ERR_CUDA, "table upload failed"); }
    *out = c;
    return 0;
}

void luxb200_ctx_destroy(luxb200_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    DevBuf* bufs[] = {&c->bsk, &c->mono, &c->kskA, &c->kskB, &c->automap, &c->loggen, &c->in0, &c->in1, &c->in2, &c->in3, &c->wa,
                      &c->wb, &c->lut, &c->ext, &c->oa, &c->ob, &c->o64a, &c->o64b, &c->sched, &c->tmpa, &c->tmpb,
                      &c->tmpc, &c->tmpd, &c->cpa, &c->cpb, &c->cdesc, &c->clsel, &c->cx1a, &c->cx1b, &c->cx2a, &c->cx2b,
                      &c->xin};
    for (DevBuf* b : bufs) b->release();
    c->wide.tw.release(); c->wide.twi.release(); c->wide.pw.release(); c->wide.automap.release();
    c->ntt.release(); c->ntt1.release();
    for (int s = 0; s < 2; ++s) {
        for (DevBuf& b : c->pipe.in[s]) b.release();
        for (DevBuf& b : c->pipe.st[s]) b.release();
        for (cudaEvent_t e : {c->pipe.e_in[s], c->pipe.e_free[s], c->pipe.e_stage[s], c->pipe.e_out[s]})
            if (e) cudaEventDestroy(e);
    }
    for (cudaEvent_t e : c->pipe.tev) cudaEventDestroy(e);
    if (c->pipe.s_in) cudaStreamDestroy(c->pipe.s_in);
    if (c->pipe.s_out) cudaStreamDestroy(c->pipe.s_out);
    for (int i = 0; i < 5; ++i)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int luxb200_ctx_info(luxb200_ctx* c, uint64_t* o) {
    if (!c || !o) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    o[0] = c->p.n; o[1] = c->p.N; o[2] = c->p.q; o[3] = c->p.Q; o[4] = c->p.qKS; o[5] = c->p.baseKS; o[6] = c->p.baseG;
    o[7] = c->p.baseR; o[8] = c->digitsG; o[9] = c->p.numAutoKeys; o[10] = 0; o[11] = c->p.psi; o[12] = c->digitsKS;
    o[13] = c->p.method; o[14] = c->digitsR; o[15] = (uint64_t)c->device;
    return 0;
}

uint64_t luxb200_bsk_len(luxb200_ctx* c) {
    if (!c) return 0;
    const uint64_t rg = (uint64_t)c->rows * 2 * c->p.N;
    if (c->p.method == LUXB200_GINX) return (uint64_t)c->p.n * 2 * rg;
    if (c->p.method == LUXB200_AP) return (uint64_t)c->p.n * c->p.baseR * c->digitsR * rg;
    return (uint64_t)c->p.n * rg + (uint64_t)(c->p.numAutoKeys + 1) * (c->digitsG - 1) * 2 * c->p.N;
}
uint64_t luxb200_ksk_len(luxb200_ctx* c) {
    if (!c) return 0;
    return (uint64_t)c->p.N * c->p.baseKS * c->digitsKS * c->p.n;
}

int luxb200_load_bsk(luxb200_ctx* c, const uint64_t* bsk, int on_device) {
    if (!c || !bsk) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    LB_TRY(lb_set_device(c->device));
    const uint64_t len = luxb200_bsk_len(c);
    if (c->wide.on && c->wide.wsize == 8) {
        // 64-bit residues in natural EVALUATION order: the canonical layout IS the device layout
        LB_TRY(c->bsk.ensure(len * 8));
        LB_CUDA(cudaMemcpyAsync(c->bsk.p, bsk, len * 8, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
        LB_CUDA(cudaStreamSynchronize(c->stream));
        c->have_bsk = true;
        return 0;
    }
    LB_TRY(c->bsk.ensure(len * 4));
    // stream the canonical u64 key through a bounded staging buffer
    const uint64_t chunk = (uint64_t)c->p.N * 4096;  // polynomials are N-aligned
    DevBuf stage;
    if (!on_device) LB_TRY(stage.ensure(std::min(len, chunk) * 8));
    for (uint64_t off = 0; off < len; off += chunk) {
        const uint64_t m = std::min(chunk, len - off);
        const u64* src = bsk + off;
        if (!on_device) {
            LB_CUDA(cudaMemcpyAsync(stage.p, bsk + off, m * 8, cudaMemcpyHostToDevice, c->stream));
            src = stage.as<u64>();
        }
        if (c->wide.on) narrow_u64_u32_kernel<<<nblk(m, 256), 256, 0, c->stream>>>(src, c->bsk.as<u32>() + off, m);
        else pack_polys_sigma_kernel<<<nblk(m, 256), 256, 0, c->stream>>>(src, c->bsk.as<u32>() + off, c->logN, m, c->quad ? 1 : 0);
        c->launches++;
    }
    LB_CUDA(cudaStreamSynchronize(c->stream));
    stage.release();
    c->have_bsk = true;
    return 0;
}

int luxb200_load_ksk(luxb200_ctx* c, const uint64_t* A, const uint64_t* B, int on_device) {
    if (!c || !A || !B) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    LB_TRY(lb_set_device(c->device));
    const uint64_t rows = (uint64_t)c->p.N * c->p.baseKS * c->digitsKS;
    LB_TRY(c->kskA.ensure(rows * c->npad * c->ks_elem));
    LB_TRY(c->kskB.ensure(rows * (c->ks_elem == 8 ? 8 : 4)));
    const uint64_t chunk_rows = 65536;
    DevBuf stage;
    if (!on_device) LB_TRY(stage.ensure(std::min(rows, chunk_rows) * c->p.n * 8));
    for (uint64_t r0 = 0; r0 < rows; r0 += chunk_rows) {
        const uint64_t m = std::min(chunk_rows, rows - r0);
        const u64* src = A + r0 * c->p.n;
        if (!on_device) {
            LB_CUDA(cudaMemcpyAsync(stage.p, src, m * c->p.n * 8, cudaMemcpyHostToDevice, c->stream));
            src = stage.as<u64>();
        }
        if (c->ks_elem == 2)
            pack_ksk_kernel<uint16_t><<<nblk(m * c->npad, 256), 256, 0, c->stream>>>(
                src, c->kskA.as<uint16_t>() + r0 * c->npad, c->p.n, c->npad, m);
        else if (c->ks_elem == 4)
            pack_ksk_kernel<u32><<<nblk(m * c->npad, 256), 256, 0, c->stream>>>(src, c->kskA.as<u32>() + r0 * c->npad,
                                                                               c->p.n, c->npad, m);
        else
            pack_ksk_kernel<u64><<<nblk(m * c->npad, 256), 256, 0, c->stream>>>(src, c->kskA.as<u64>() + r0 * c->npad,
                                                                               c->p.n, c->npad, m);
        c->launches++;
    }
    {
        DevBuf sb;
        const u64* src = B;
        if (!on_device) {
            LB_TRY(sb.ensure(rows * 8));
            LB_CUDA(cudaMemcpyAsync(sb.p, B, rows * 8, cudaMemcpyHostToDevice, c->stream));
            src = sb.as<u64>();
        }
        if (c->ks_elem == 8) copy_u64_kernel<<<nblk(rows, 256), 256, 0, c->stream>>>(src, c->kskB.as<u64>(), rows);
        else narrow_u64_u32_kernel<<<nblk(rows, 256), 256, 0, c->stream>>>(src, c->kskB.as<u32>(), rows);
        c->launches++;
        LB_CUDA(cudaStreamSynchronize(c->stream));
        sb.release();
    }
    stage.release();
    c->have_ksk = true;
    return 0;
}

int luxb200_synchronize(luxb200_ctx* c) {
    if (!c) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    LB_TRY(lb_set_device(c->device));
    LB_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}


// ---- pinned host staging (cudaHostAlloc): host-side packers write ciphertext batches straight into page-locked memory so
// the copy-in / copy-out streams of the batch pipeline run at PCIe speed and truly overlap the kernels
int luxb200_host_alloc(uint64_t bytes, void** out) {
    if (!out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    int cnt = 0;
    if (cudaGetDeviceCount(&cnt) != cudaSuccess || cnt == 0)
        return lb_fail(LUXB200_ERR_NO_DEVICE, "no CUDA device available; libluxb200 has no CPU fallback");
    if (cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return lb_fail(LUXB200_ERR_ALLOC, "cudaHostAlloc(%llu) failed", (unsigned long long)bytes);
    }
    return 0;
}
int luxb200_host_free(void* p) {
    if (!p) return 0;
    LB_CUDA(cudaFreeHost(p));
    return 0;
}

// ---- one-time key broadcast inside one process: ctxs[0] holds the keys (luxb200_load_bsk / _ksk); every other context
// (same parameters, another GPU of the box) receives the PACKED device images — u32 sigma-ordered BSK, u16/u32 KSK rows,
// u32 KSK B — by peer-to-peer copies over NVLink/NVSwitch (cudaMemcpyPeerAsync; staged through the host by the driver
// when two devices have no peer path).  Doubling tree: after round r, 2^r contexts hold the keys and each sends to one
// more, so all egress links work and n devices take ceil(log2 n) rounds.  No collective exists on the per-gate path.
int luxb200_broadcast_keys(luxb200_ctx* const* ctxs, int n) {
    if (!ctxs || n <= 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    luxb200_ctx* r = ctxs[0];
    if (!r) return lb_fail(LUXB200_ERR_INVALID_ARG, "null context");
    if (!r->have_bsk && !r->have_ksk) return lb_fail(LUXB200_ERR_NO_KEYS, "luxb200_broadcast_keys: the root context holds no keys");
    for (int i = 1; i < n; ++i) {
        luxb200_ctx* c = ctxs[i];
        if (!c) return lb_fail(LUXB200_ERR_INVALID_ARG, "null context");
        if (memcmp(&c->p, &r->p, sizeof(luxb200_params)) != 0)
            return lb_fail(LUXB200_ERR_INVALID_ARG, "luxb200_broadcast_keys: context %d has different parameters", i);
        for (int j = 0; j < i; ++j)
            if (ctxs[j] == c) return lb_fail(LUXB200_ERR_INVALID_ARG, "luxb200_broadcast_keys: context %d listed twice", i);
    }
    // whichever of the two keys the root holds is sent (a backend may bind a refresh key before a switching key exists)
    const size_t bsk_bytes = r->have_bsk ? luxb200_bsk_len(r) * ((r->wide.on && r->wide.wsize == 8) ? 8 : 4) : 0;
    const size_t rowsK = (size_t)r->p.N * r->p.baseKS * r->digitsKS;
    const size_t ka_bytes = r->have_ksk ? rowsK * r->npad * r->ks_elem : 0, kb_bytes = r->have_ksk ? rowsK * (r->ks_elem == 8 ? 8 : 4) : 0;
    for (int i = 1; i < n; ++i) {
        luxb200_ctx* c = ctxs[i];
        LB_TRY(lb_set_device(c->device));
        if (bsk_bytes) LB_TRY(c->bsk.ensure(bsk_bytes));
        if (ka_bytes) { LB_TRY(c->kskA.ensure(ka_bytes)); LB_TRY(c->kskB.ensure(kb_bytes)); }
        if (c->device != r->device) {
            int can = 0;
            cudaDeviceCanAccessPeer(&can, c->device, r->device);
            if (can && cudaDeviceEnablePeerAccess(r->device, 0) == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        }
    }
    for (int have = 1; have < n; have *= 2) {
        const int m = std::min(have, n - have);
        for (int k = 0; k < m; ++k) {   // ctxs[k] -> ctxs[have + k], on the receiver's stream
            luxb200_ctx *src = ctxs[k], *dst = ctxs[have + k];
            LB_TRY(lb_set_device(dst->device));
            if (bsk_bytes) LB_CUDA(cudaMemcpyPeerAsync(dst->bsk.p, dst->device, src->bsk.p, src->device, bsk_bytes, dst->stream));
            if (ka_bytes) {
                LB_CUDA(cudaMemcpyPeerAsync(dst->kskA.p, dst->device, src->kskA.p, src->device, ka_bytes, dst->stream));
                LB_CUDA(cudaMemcpyPeerAsync(dst->kskB.p, dst->device, src->kskB.p, src->device, kb_bytes, dst->stream));
            }
        }
        for (int k = 0; k < m; ++k) {
            luxb200_ctx* dst = ctxs[have + k];
            LB_TRY(lb_set_device(dst->device));
            LB_CUDA(cudaStreamSynchronize(dst->stream));
            dst->have_bsk = r->have_bsk;
            dst->have_ksk = r->have_ksk;
        }
    }
    return 0;
}

}  // extern "C"

#include "api_ops.inc"
