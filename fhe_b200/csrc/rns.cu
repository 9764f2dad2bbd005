// rns.cu — the RNS basis-conversion steps either side of the N = 2^16 NTT in CKKS/BGV hybrid key switching:
//   DCRTPolyImpl::ApproxSwitchCRTBasis   src/core/include/lattice/hal/default/dcrtpoly-impl.h:910-957
//   DCRTPolyImpl::ApproxModUp            :960-986    (KeySwitchHYBRID::EvalKeySwitchPrecomputeCore, keyswitch-hybrid.cpp:328-412)
//   DCRTPolyImpl::ApproxModDown          :988-1030   (KeySwitchHYBRID::EvalFastKeySwitchCore, keyswitch-hybrid.cpp:414-436; t = 0, CKKS)
// out_j[k] = ( sum_i [x_i[k] * (Q/q_i)^-1]_{q_i} * [(Q/q_i)]_{p_j} ) mod p_j — the reference accumulates the products in 128 bits
// and reduces once (BarrettUint128ModUint64); every stored value is the canonical residue, so any exact evaluation of the same
// integer sum is bit-identical.  Tables (QHatInvModq, QHatModp, PInvModq) are derived from the moduli on the host exactly as
// CryptoParametersRNS::PrecomputeCRTTables defines them.  HBM-resident [count][limbs][N] uint64, one thread per coefficient,
// limb tables in shared memory; the NTTs in between are luxb200_ntt_batch_dev (ntt64_tma.inc for N >= 2^14).
#include <algorithm>

#include "api_common.h"
#include "modarith.cuh"

typedef unsigned __int128 u128;

namespace {

inline unsigned nblk(u64 total, unsigned per) { return (unsigned)((total + per - 1) / per); }

struct RnsDev {
    // per source limb i: modulus, (Q/q_i)^-1 mod q_i and its Shoup companion; per target limb j: modulus, floor(2^64 / p_j),
    // 2^64 mod p_j and its Shoup companion; hat [sizeQ][sizeP] = (Q/q_i) mod p_j
    const u64 *q, *qhatinv, *qhatinv_s, *p, *pmu, *pr, *prs, *hat;
    u32 sizeQ, sizeP, N, pad;
};

// x [count][in_stride limbs][N] (the first sizeQ limbs are read), out [count][out_stride limbs][N] (limbs out_off .. out_off+sizeP-1)
__global__ void __launch_bounds__(256) rns_switch_kernel(const RnsDev D, const u64* __restrict__ x, u32 in_stride, u32 in_off,
                                                         u64* __restrict__ out, u32 out_stride, u32 out_off, u64 count) {
    extern __shared__ u64 rns_sm[];
    u64* hat = rns_sm;                                   // [sizeQ][sizeP]
    u64* xh  = rns_sm + (size_t)D.sizeQ * D.sizeP;       // [sizeQ][256]
    for (u32 t = threadIdx.x; t < D.sizeQ * D.sizeP; t += blockDim.x) hat[t] = D.hat[t];
    const u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = gid < count * D.N;
    const u64 ct = live ? gid / D.N : 0;
    const u32 k = (u32)(gid % D.N);
    if (live) {
        for (u32 i = 0; i < D.sizeQ; ++i) {
            const u64 v = x[(ct * in_stride + in_off + i) * D.N + k];
            const u64 qi = D.q[i];
            u64 r = v * D.qhatinv[i] - __umul64hi(v, D.qhatinv_s[i]) * qi;   // Shoup: [0, 2 q_i)
            r = r >= qi ? r - qi : r;
            xh[(size_t)i * 256 + threadIdx.x] = r;
        }
    }
    __syncthreads();
    if (!live) return;
    for (u32 j = 0; j < D.sizeP; ++j) {
        u128 s = 0;
        for (u32 i = 0; i < D.sizeQ; ++i) s += (u128)xh[(size_t)i * 256 + threadIdx.x] * hat[(size_t)i * D.sizeP + j];
        const u64 pj = D.p[j], lo = (u64)s, hi = (u64)(s >> 64);
        const u64 t = hi * D.pr[j] - __umul64hi(hi, D.prs[j]) * pj;      // hi * 2^64 mod p_j, in [0, 2 p_j)
        const u64 l = lo - __umul64hi(lo, D.pmu[j]) * pj;                // lo mod p_j, in [0, 2 p_j)
        u64 r = t + l;                                                   // < 4 p_j < 2^64 (p_j < 2^62)
        r = r >= 2 * pj ? r - 2 * pj : r;
        r = r >= pj ? r - pj : r;
        out[(ct * out_stride + out_off + j) * D.N + k] = r;
    }
}

// ApproxModDown tail: out_i = (x_i - s_i) * PInvModq_i mod q_i   (dcrtpoly-impl.h:1021-1028), EVALUATION format
__global__ void rns_moddown_tail_kernel(const u64* __restrict__ x, u32 x_stride, const u64* __restrict__ s, u64* __restrict__ out,
                                        const u64* __restrict__ q, const u64* __restrict__ pinv, const u64* __restrict__ pinv_s,
                                        u32 sizeQ, u32 N, u64 count) {
    const u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= count * sizeQ * N) return;
    const u32 k = (u32)(gid % N);
    const u32 i = (u32)((gid / N) % sizeQ);
    const u64 ct = gid / ((u64)N * sizeQ);
    const u64 qi = q[i];
    const u64 a = x[(ct * x_stride + i) * N + k], b = s[gid];
    const u64 d = a >= b ? a - b : a + qi - b;
    u64 r = d * pinv[i] - __umul64hi(d, pinv_s[i]) * qi;
    out[gid] = r >= qi ? r - qi : r;
}

// 128-bit sum -> canonical residue mod m (m < 2^62): the high word folded with a Shoup multiplication by r = 2^64 mod m, the low
// word with a Barrett step (mu = floor(2^64 / m)), both in [0, 2m)
__device__ __forceinline__ u64 rns_reduce128(u128 s, u64 m, u64 mu, u64 r, u64 rs) {
    const u64 lo = (u64)s, hi = (u64)(s >> 64);
    const u64 t = hi * r - __umul64hi(hi, rs) * m;
    const u64 l = lo - __umul64hi(lo, mu) * m;
    u64 v = t + l;
    v = v >= 2 * m ? v - 2 * m : v;
    return v >= m ? v - m : v;
}

// KeySwitchHYBRID::EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:438-481): out_c[i][k] = sum_j digits_j[i][k] * evk_c[j][idx(i)][k]
// mod m_i over the basis Ql u P; one thread per (ciphertext, limb, coefficient), exact 128-bit sums reduced once per four terms
// (any exact evaluation gives the reference's canonical residues).  HBM-bound: 24 bytes read per digit, 16 written.
struct RnsInner {
    const u64 *qm, *qmu, *qr, *qrs;   // Q moduli and their reduction constants
    const u64 *pm, *pmu, *pr, *prs;   // P moduli
    u32 sizeQ, sizeP, sizeQl, nd, N, pad;
};
__global__ void __launch_bounds__(256) rns_inner_kernel(const RnsInner D, const u64* __restrict__ digits, const u64* __restrict__ kb,
                                                        const u64* __restrict__ ka, u64* __restrict__ out, u64 count) {
    const u32 L = D.sizeQl + D.sizeP, LK = D.sizeQ + D.sizeP;
    const u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= count * L * D.N) return;
    const u32 k = (u32)(gid % D.N);
    const u32 i = (u32)((gid / D.N) % L);
    const u64 c = gid / ((u64)D.N * L);
    const bool isq = i < D.sizeQl;
    const u32 li = isq ? i : i - D.sizeQl, idx = isq ? i : D.sizeQ + li;
    const u64 m = isq ? D.qm[li] : D.pm[li], mu = isq ? D.qmu[li] : D.pmu[li], r = isq ? D.qr[li] : D.pr[li], rs = isq ? D.qrs[li] : D.prs[li];
    u128 s0 = 0, s1 = 0;
    for (u32 j = 0; j < D.nd; ++j) {
        const u64 d = digits[((c * D.nd + j) * L + i) * D.N + k];
        s0 += (u128)d * kb[((u64)j * LK + idx) * D.N + k];
        s1 += (u128)d * ka[((u64)j * LK + idx) * D.N + k];
        if ((j & 3) == 3) { s0 = rns_reduce128(s0, m, mu, r, rs); s1 = rns_reduce128(s1, m, mu, r, rs); }   // 4 m^2 + m < 2^128
    }
    out[((c * 2 + 0) * L + i) * D.N + k] = rns_reduce128(s0, m, mu, r, rs);
    out[((c * 2 + 1) * L + i) * D.N + k] = rns_reduce128(s1, m, mu, r, rs);
}

// data[c][l][k] <- data[c][l][k] * w_l mod m_l (canonical in, canonical out): the BGV factors of ApproxModDown, t^-1 mod p_j on the
// P part before the basis switch and t mod q_i on the switched part (dcrtpoly-impl.h:1003-1005, :1021-1023)
__global__ void rns_scale_limbs_kernel(u64* __restrict__ data, const u64* __restrict__ mod, const u64* __restrict__ w,
                                       const u64* __restrict__ ws, u32 limbs, u32 N, u64 count) {
    const u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= count * limbs * N) return;
    const u32 l = (u32)((gid / N) % limbs);
    const u64 m = mod[l], v = data[gid];
    const u64 r = v * w[l] - __umul64hi(v, ws[l]) * m;
    data[gid] = r >= m ? r - m : r;
}

__global__ void rns_copy_limbs_kernel(const u64* __restrict__ in, u32 in_stride, u32 in_off, u64* __restrict__ out, u32 out_stride,
                                      u32 out_off, u32 limbs, u32 N, u64 count) {
    const u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= count * limbs * N) return;
    const u32 k = (u32)(gid % N);
    const u32 l = (u32)((gid / N) % limbs);
    const u64 ct = gid / ((u64)N * limbs);
    out[(ct * out_stride + out_off + l) * N + k] = in[(ct * in_stride + in_off + l) * N + k];
}

struct Tables {
    std::vector<u64> q, p, qhatinv, qhatinv_s, pmu, pr, prs, hat;
};

// QHatInvModq_i = ((Q / q_i) mod q_i)^-1, QHatModp[i][j] = (Q / q_i) mod p_j
void make_tables(const u64* q, u32 sizeQ, const u64* p, u32 sizeP, Tables& T) {
    using namespace lb;
    T.q.assign(q, q + sizeQ); T.p.assign(p, p + sizeP);
    T.qhatinv.resize(sizeQ); T.qhatinv_s.resize(sizeQ); T.hat.resize((size_t)sizeQ * sizeP);
    T.pmu.resize(sizeP); T.pr.resize(sizeP); T.prs.resize(sizeP);
    for (u32 i = 0; i < sizeQ; ++i) {
        u64 h = 1;
        for (u32 k = 0; k < sizeQ; ++k)
            if (k != i) h = mulmod(h, q[k] % q[i], q[i]);
        T.qhatinv[i] = invmod_prime(h, q[i]);
        T.qhatinv_s[i] = shoup64(T.qhatinv[i], q[i]);
        for (u32 j = 0; j < sizeP; ++j) {
            u64 g = 1;
            for (u32 k = 0; k < sizeQ; ++k)
                if (k != i) g = mulmod(g, q[k] % p[j], p[j]);
            T.hat[(size_t)i * sizeP + j] = g;
        }
    }
    for (u32 j = 0; j < sizeP; ++j) {
        T.pmu[j] = (u64)((((u128)1) << 64) / p[j]);
        T.pr[j] = (u64)((((u128)1) << 64) % p[j]);
        T.prs[j] = shoup64(T.pr[j], p[j]);
    }
}

}  // namespace

struct luxb200_rns_plan {
    uint32_t N = 0, sizeQ = 0, sizeP = 0;
    int device = 0;
    Tables up, down;            // Q -> P (mod up) and P -> Q (mod down)
    std::vector<u64> pinv, pinv_s;   // P^-1 mod q_i
    DevBuf dev_up, dev_down, dev_tail, tmpP, tmpQ, stage_in, stage_out, dev_t;
    uint64_t t_cached = 0;            // plaintext modulus whose BGV factors are resident in dev_t
    std::vector<u64> t_host;          // [p | tinv | tinv_s | q | tq | tq_s]
    RnsDev d_up{}, d_down{};
    const u64 *d_q = nullptr, *d_pinv = nullptr, *d_pinv_s = nullptr;
    luxb200_ntt_plan *nttQ = nullptr, *nttP = nullptr;
    cudaStream_t stream = nullptr;
};

struct luxb200_hks_plan {
    uint32_t N = 0, sizeQ = 0, sizeP = 0, alpha = 0, sizeQl = 0, nd = 0;
    int device = 0;
    std::vector<luxb200_rns_plan*> parts;              // per digit: source = the part's limbs, target = every other limb of Ql u P
    luxb200_rns_plan *full = nullptr, *down = nullptr; // (Q, P): inner-product constants; (Ql, P): ApproxModDown
    DevBuf cin, part_in, part_out, digits, tilda, key, outb;
};

namespace {

int upload_tables(const Tables& T, DevBuf& buf, RnsDev& D, u32 N) {
    const u32 sq = (u32)T.q.size(), sp = (u32)T.p.size();
    std::vector<u64> flat;
    auto push = [&](const std::vector<u64>& v) { size_t o = flat.size(); flat.insert(flat.end(), v.begin(), v.end()); return o; };
    const size_t oq = push(T.q), oi = push(T.qhatinv), ois = push(T.qhatinv_s), op = push(T.p), omu = push(T.pmu), opr = push(T.pr),
                 oprs = push(T.prs), oh = push(T.hat);
    LB_TRY(buf.ensure(flat.size() * 8));
    LB_CUDA(cudaMemcpy(buf.p, flat.data(), flat.size() * 8, cudaMemcpyHostToDevice));
    const u64* b = buf.as<u64>();
    D = RnsDev{b + oq, b + oi, b + ois, b + op, b + omu, b + opr, b + oprs, b + oh, sq, sp, N, 0};
    return 0;
}

int launch_switch(luxb200_rns_plan* pl, const RnsDev& D, const u64* x, u32 in_stride, u32 in_off, u64* out, u32 out_stride,
                  u32 out_off, u64 count) {
    const size_t sm = ((size_t)D.sizeQ * D.sizeP + (size_t)D.sizeQ * 256) * 8;
    if (sm > 48 * 1024) LB_CUDA(cudaFuncSetAttribute(rns_switch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    rns_switch_kernel<<<nblk(count * D.N, 256), 256, sm, pl->stream>>>(D, x, in_stride, in_off, out, out_stride, out_off, count);
    LB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int luxb200_rns_plan_create(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, int device,
                            luxb200_rns_plan** out) {
    if (!q || !p || !out || sizeQ == 0 || sizeP == 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    if (sizeQ > 64 || sizeP > 64) return lb_fail(LUXB200_ERR_UNSUPPORTED, "at most 64 limbs per basis");
    for (uint32_t i = 0; i < sizeQ + sizeP; ++i) {
        const uint64_t m = i < sizeQ ? q[i] : p[i - sizeQ];
        if (m >= (1ull << 62) || !lb::is_prime(m)) return lb_fail(LUXB200_ERR_INVALID_ARG, "moduli must be primes below 2^62");
        for (uint32_t k = 0; k < i; ++k)
            if ((k < sizeQ ? q[k] : p[k - sizeQ]) == m) return lb_fail(LUXB200_ERR_INVALID_ARG, "moduli must be pairwise distinct");
    }
    LB_TRY(lb_set_device(device));
    auto* pl = new luxb200_rns_plan();
    pl->N = N; pl->sizeQ = sizeQ; pl->sizeP = sizeP; pl->device = device;
    make_tables(q, sizeQ, p, sizeP, pl->up);
    make_tables(p, sizeP, q, sizeQ, pl->down);
    pl->pinv.resize(sizeQ); pl->pinv_s.resize(sizeQ);
    for (uint32_t i = 0; i < sizeQ; ++i) {
        uint64_t P = 1;
        for (uint32_t j = 0; j < sizeP; ++j) P = lb::mulmod(P, p[j] % q[i], q[i]);
        pl->pinv[i] = lb::invmod_prime(P, q[i]);
        pl->pinv_s[i] = lb::shoup64(pl->pinv[i], q[i]);
    }
    int rc = upload_tables(pl->up, pl->dev_up, pl->d_up, N);
    if (!rc) rc = upload_tables(pl->down, pl->dev_down, pl->d_down, N);
    if (!rc) {
        std::vector<u64> flat(q, q + sizeQ);
        flat.insert(flat.end(), pl->pinv.begin(), pl->pinv.end());
        flat.insert(flat.end(), pl->pinv_s.begin(), pl->pinv_s.end());
        rc = pl->dev_tail.ensure(flat.size() * 8);
        if (!rc && cudaMemcpy(pl->dev_tail.p, flat.data(), flat.size() * 8, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = lb_fail(LUXB200_ERR_CUDA, "table upload failed");
        pl->d_q = pl->dev_tail.as<u64>(); pl->d_pinv = pl->d_q + sizeQ; pl->d_pinv_s = pl->d_pinv + sizeQ;
    }
    // the NTTs between the steps: only when N is a transform size (the raw basis switch works for any N)
    bool ntt_ok = N >= 256 && N <= 131072 && !(N & (N - 1));
    for (uint32_t i = 0; i < sizeQ + sizeP && ntt_ok; ++i) ntt_ok = ((i < sizeQ ? q[i] : p[i - sizeQ]) - 1) % (2ull * N) == 0;
    if (!rc && ntt_ok) {
        rc = luxb200_ntt_plan_create(N, sizeQ, q, nullptr, device, &pl->nttQ);
        if (!rc) rc = luxb200_ntt_plan_create(N, sizeP, p, nullptr, device, &pl->nttP);
    }
    if (!rc && cudaStreamCreateWithFlags(&pl->stream, cudaStreamNonBlocking) != cudaSuccess) rc = lb_fail(LUXB200_ERR_CUDA, "stream creation failed");
    if (!rc && cudaDeviceSynchronize() != cudaSuccess) rc = lb_fail(LUXB200_ERR_CUDA, "table upload failed");   // the uploads above have landed
    if (rc) { luxb200_rns_plan_destroy(pl); return rc; }
    *out = pl;
    return 0;
}

void luxb200_rns_plan_destroy(luxb200_rns_plan* pl) {
    if (!pl) return;
    cudaSetDevice(pl->device);
    if (pl->nttQ) luxb200_ntt_plan_destroy(pl->nttQ);
    if (pl->nttP) luxb200_ntt_plan_destroy(pl->nttP);
    if (pl->stream) cudaStreamDestroy(pl->stream);
    delete pl;
}

// DCRTPolyImpl::ApproxSwitchCRTBasis: x [count][sizeQ][N] (COEFFICIENT format, residues mod q_i) -> out [count][sizeP][N];
// direction 0: Q -> P, 1: P -> Q (then x has sizeP limbs and out sizeQ).  Host buffers.
int luxb200_approx_switch_crt_basis(luxb200_rns_plan* pl, const uint64_t* x, uint64_t* out, uint64_t count, int direction) {
    if (!pl || !x || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(pl->device));
    const RnsDev& D = direction ? pl->d_down : pl->d_up;
    const uint64_t in_el = (uint64_t)D.sizeQ * pl->N, out_el = (uint64_t)D.sizeP * pl->N;
    const uint64_t chunk = std::max<uint64_t>(1, (256ull << 20) / ((in_el + out_el) * 8));
    for (uint64_t off = 0; off < count; off += chunk) {
        const uint64_t m = std::min(chunk, count - off);
        LB_TRY(pl->stage_in.ensure(m * in_el * 8)); LB_TRY(pl->stage_out.ensure(m * out_el * 8));
        LB_CUDA(cudaMemcpyAsync(pl->stage_in.p, x + off * in_el, m * in_el * 8, cudaMemcpyHostToDevice, pl->stream));
        LB_TRY(launch_switch(pl, D, pl->stage_in.as<u64>(), D.sizeQ, 0, pl->stage_out.as<u64>(), D.sizeP, 0, m));
        LB_CUDA(cudaMemcpyAsync(out + off * out_el, pl->stage_out.p, m * out_el * 8, cudaMemcpyDeviceToHost, pl->stream));
        LB_CUDA(cudaStreamSynchronize(pl->stream));
    }
    return 0;
}

// Device-resident composites, EVALUATION format in and out (data on the plan's device, asynchronous on the plan's stream order:
// each step is followed by a stream synchronisation because the NTT plans own their streams).
// ApproxModUp: x [count][sizeQ][N] -> out [count][sizeQ + sizeP][N] = the Q limbs unchanged, then the P limbs.
int luxb200_approx_mod_up_dev(luxb200_rns_plan* pl, const void* x_dev, void* out_dev, uint64_t count) {
    if (!pl || !x_dev || !out_dev) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (!pl->nttQ) return lb_fail(LUXB200_ERR_UNSUPPORTED, "this plan has no NTT (N or the moduli do not admit one)");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(pl->device));
    const uint32_t N = pl->N, sq = pl->sizeQ, sp = pl->sizeP;
    const u64* x = reinterpret_cast<const u64*>(x_dev);
    u64* out = reinterpret_cast<u64*>(out_dev);
    LB_TRY(pl->tmpQ.ensure(count * sq * N * 8)); LB_TRY(pl->tmpP.ensure(count * sp * N * 8));
    rns_copy_limbs_kernel<<<nblk(count * sq * N, 256), 256, 0, pl->stream>>>(x, sq, 0, pl->tmpQ.as<u64>(), sq, 0, sq, N, count);
    rns_copy_limbs_kernel<<<nblk(count * sq * N, 256), 256, 0, pl->stream>>>(x, sq, 0, out, sq + sp, 0, sq, N, count);
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    LB_TRY(luxb200_ntt_batch_dev(pl->nttQ, pl->tmpQ.p, count, 1, 8));          // SetFormat(COEFFICIENT)
    LB_TRY(luxb200_ntt_plan_sync(pl->nttQ));
    LB_TRY(launch_switch(pl, pl->d_up, pl->tmpQ.as<u64>(), sq, 0, pl->tmpP.as<u64>(), sp, 0, count));
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    LB_TRY(luxb200_ntt_batch_dev(pl->nttP, pl->tmpP.p, count, 0, 8));          // the new limbs to EVALUATION
    LB_TRY(luxb200_ntt_plan_sync(pl->nttP));
    rns_copy_limbs_kernel<<<nblk(count * sp * N, 256), 256, 0, pl->stream>>>(pl->tmpP.as<u64>(), sp, 0, out, sq + sp, sq, sp, N, count);
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    return 0;
}

// ApproxModDown (t = 0, CKKS): x [count][sizeQ + sizeP][N] EVALUATION -> out [count][sizeQ][N] EVALUATION
int luxb200_approx_mod_down_dev(luxb200_rns_plan* pl, const void* x_dev, void* out_dev, uint64_t count) {
    return luxb200_approx_mod_down_bgv_dev(pl, x_dev, out_dev, count, 0);
}
// ApproxModDown with plaintext modulus t (BGV): the P part times t^-1 mod p_j before the switch, the switched part times t mod q_i
int luxb200_approx_mod_down_bgv_dev(luxb200_rns_plan* pl, const void* x_dev, void* out_dev, uint64_t count, uint64_t t) {
    if (!pl || !x_dev || !out_dev) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (!pl->nttQ) return lb_fail(LUXB200_ERR_UNSUPPORTED, "this plan has no NTT (N or the moduli do not admit one)");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(pl->device));
    const uint32_t N = pl->N, sq = pl->sizeQ, sp = pl->sizeP;
    const u64* x = reinterpret_cast<const u64*>(x_dev);
    u64* out = reinterpret_cast<u64*>(out_dev);
    LB_TRY(pl->tmpQ.ensure(count * sq * N * 8)); LB_TRY(pl->tmpP.ensure(count * sp * N * 8));
    const u64 *d_tp = nullptr, *d_tq = nullptr;
    if (t) {
        if (pl->t_cached != t) {
            std::vector<u64>& h = pl->t_host;
            h.assign(3 * (size_t)(sq + sp), 0);
            for (uint32_t j = 0; j < sp; ++j) {
                const u64 m = pl->up.p[j];
                if (t % m == 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "t has no inverse modulo an auxiliary prime");
                const u64 inv = lb::invmod_prime(t % m, m);
                h[j] = m; h[sp + j] = inv; h[2 * sp + j] = lb::shoup64(inv, m);
            }
            for (uint32_t i = 0; i < sq; ++i) {
                const u64 m = pl->up.q[i];
                h[3 * sp + i] = m; h[3 * sp + sq + i] = t % m; h[3 * sp + 2 * sq + i] = lb::shoup64(t % m, m);
            }
            LB_TRY(pl->dev_t.ensure(h.size() * 8));
            LB_CUDA(cudaMemcpyAsync(pl->dev_t.p, h.data(), h.size() * 8, cudaMemcpyHostToDevice, pl->stream));
            LB_CUDA(cudaStreamSynchronize(pl->stream));
            pl->t_cached = t;
        }
        d_tp = pl->dev_t.as<u64>(); d_tq = d_tp + 3 * sp;
    }
    rns_copy_limbs_kernel<<<nblk(count * sp * N, 256), 256, 0, pl->stream>>>(x, sq + sp, sq, pl->tmpP.as<u64>(), sp, 0, sp, N, count);
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    LB_TRY(luxb200_ntt_batch_dev(pl->nttP, pl->tmpP.p, count, 1, 8));
    LB_TRY(luxb200_ntt_plan_sync(pl->nttP));
    if (t) rns_scale_limbs_kernel<<<nblk(count * sp * N, 256), 256, 0, pl->stream>>>(pl->tmpP.as<u64>(), d_tp, d_tp + sp, d_tp + 2 * sp, sp, N, count);
    LB_TRY(launch_switch(pl, pl->d_down, pl->tmpP.as<u64>(), sp, 0, pl->tmpQ.as<u64>(), sq, 0, count));
    if (t) rns_scale_limbs_kernel<<<nblk(count * sq * N, 256), 256, 0, pl->stream>>>(pl->tmpQ.as<u64>(), d_tq, d_tq + sq, d_tq + 2 * sq, sq, N, count);
    LB_CUDA(cudaGetLastError());
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    LB_TRY(luxb200_ntt_batch_dev(pl->nttQ, pl->tmpQ.p, count, 0, 8));
    LB_TRY(luxb200_ntt_plan_sync(pl->nttQ));
    rns_moddown_tail_kernel<<<nblk(count * sq * N, 256), 256, 0, pl->stream>>>(x, sq + sp, pl->tmpQ.as<u64>(), out, pl->d_q, pl->d_pinv,
                                                                             pl->d_pinv_s, sq, N, count);
    LB_CUDA(cudaGetLastError());
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    return 0;
}

// ---- allocation-free variant of luxb200_approx_switch_crt_basis ------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops, harnesses that forbid it).  The device staging the basis-switch kernel reads its input
// from and writes its result to lives in a file-scope array instead of the plan's growable DevBufs: every GPU that loads the
// library holds LUXB200_STATIC_SCRATCH_BYTES of it until exit.  All calls on one device share the array and it keeps the last
// call's contents; nothing has to start at zero (the input part is overwritten by the H2D copy and every output word is
// written by the kernel before the D2H copy reads it), but two calls that overlap on one GPU race on it — serialise them.
#define LB_STATIC_SCRATCH_WORDS (LUXB200_STATIC_SCRATCH_BYTES / 8)
static __device__ __align__(256) unsigned long long lb_static_scratch[LB_STATIC_SCRATCH_WORDS];

namespace {
// device address of lb_static_scratch per device: in host code the array's name is only a handle for the runtime, so the
// address is asked for once per device, with that device current
std::mutex g_scratch_mu;
u64* g_scratch_ptr[64] = {};

int static_scratch(int device, u64** out) {
    if (device < 0 || device >= 64) return lb_fail(LUXB200_ERR_INVALID_ARG, "device index out of range");
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[device]) {
        LB_TRY(lb_set_device(device));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, lb_static_scratch);
        if (e != cudaSuccess || !p) return lb_fail(LUXB200_ERR_CUDA, "cudaGetSymbolAddress(lb_static_scratch): %s", cudaGetErrorString(e));
        g_scratch_ptr[device] = reinterpret_cast<u64*>(p);
    }
    *out = g_scratch_ptr[device];
    return 0;
}
}  // namespace

int luxb200_approx_switch_crt_basis_static(luxb200_rns_plan* pl, const uint64_t* x, uint64_t* out, uint64_t count, int direction) {
    if (!pl || !x || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (count == 0) return 0;
    const RnsDev& D = direction ? pl->d_down : pl->d_up;
    const uint64_t in_el = (uint64_t)D.sizeQ * pl->N, out_el = (uint64_t)D.sizeP * pl->N;
    // input staging is rounded up to 32 words so that the output staging stays 256-byte aligned
    if (count > LB_STATIC_SCRATCH_WORDS / (in_el + out_el) ||
        ((count * in_el + 31) & ~31ull) + count * out_el > LB_STATIC_SCRATCH_WORDS)
        return lb_fail(LUXB200_ERR_INVALID_ARG,
                       "luxb200_approx_switch_crt_basis_static: count * (sizeQ + sizeP) * N * 8 bytes exceeds the static scratch of "
                       "%llu bytes; split the batch or use luxb200_approx_switch_crt_basis",
                       (unsigned long long)LUXB200_STATIC_SCRATCH_BYTES);
    LB_TRY(lb_set_device(pl->device));
    u64* scratch = nullptr;
    LB_TRY(static_scratch(pl->device, &scratch));
    u64* sin = scratch;
    u64* sout = scratch + ((count * in_el + 31) & ~31ull);
    LB_CUDA(cudaMemcpyAsync(sin, x, count * in_el * 8, cudaMemcpyHostToDevice, pl->stream));
    LB_TRY(launch_switch(pl, D, sin, D.sizeQ, 0, sout, D.sizeP, 0, count));
    LB_CUDA(cudaMemcpyAsync(out, sout, count * out_el * 8, cudaMemcpyDeviceToHost, pl->stream));
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    return 0;
}

// EvalFastKeySwitchCoreExt: digits [count][nd][sizeQl + sizeP][N], evk_b / evk_a [nd][sizeQ + sizeP][N] (the key on the full
// basis), out [count][2][sizeQl + sizeP][N]; everything EVALUATION format on the plan's device
int luxb200_rns_keyswitch_inner_dev(luxb200_rns_plan* pl, const void* digits_dev, const void* evk_b_dev, const void* evk_a_dev,
                                    void* out_dev, uint64_t count, uint32_t nd, uint32_t sizeQl) {
    if (!pl || !digits_dev || !evk_b_dev || !evk_a_dev || !out_dev) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (nd == 0 || sizeQl == 0 || sizeQl > pl->sizeQ) return lb_fail(LUXB200_ERR_INVALID_ARG, "need 1 <= sizeQl <= sizeQ and at least one digit");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(pl->device));
    RnsInner D;
    D.qm = pl->d_down.p; D.qmu = pl->d_down.pmu; D.qr = pl->d_down.pr; D.qrs = pl->d_down.prs;   // the P -> Q tables target Q
    D.pm = pl->d_up.p;   D.pmu = pl->d_up.pmu;   D.pr = pl->d_up.pr;   D.prs = pl->d_up.prs;
    D.sizeQ = pl->sizeQ; D.sizeP = pl->sizeP; D.sizeQl = sizeQl; D.nd = nd; D.N = pl->N; D.pad = 0;
    const u64 total = count * (sizeQl + pl->sizeP) * pl->N;
    rns_inner_kernel<<<nblk(total, 256), 256, 0, pl->stream>>>(D, reinterpret_cast<const u64*>(digits_dev), reinterpret_cast<const u64*>(evk_b_dev),
                                                             reinterpret_cast<const u64*>(evk_a_dev), reinterpret_cast<u64*>(out_dev), count);
    LB_CUDA(cudaGetLastError());
    LB_CUDA(cudaStreamSynchronize(pl->stream));
    return 0;
}
// host buffers (the key is uploaded with every call: callers that keep it resident use the _dev form)
int luxb200_rns_keyswitch_inner(luxb200_rns_plan* pl, const uint64_t* digits, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out,
                                uint64_t count, uint32_t nd, uint32_t sizeQl) {
    if (!pl || !digits || !evk_b || !evk_a || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (nd == 0 || sizeQl == 0 || sizeQl > pl->sizeQ) return lb_fail(LUXB200_ERR_INVALID_ARG, "need 1 <= sizeQl <= sizeQ and at least one digit");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(pl->device));
    const uint64_t L = sizeQl + pl->sizeP, LK = pl->sizeQ + pl->sizeP, N = pl->N;
    const uint64_t key_el = (uint64_t)nd * LK * N, in_el = (uint64_t)nd * L * N, out_el = 2 * L * N;
    LB_TRY(pl->tmpQ.ensure(2 * key_el * 8));
    LB_CUDA(cudaMemcpyAsync(pl->tmpQ.p, evk_b, key_el * 8, cudaMemcpyHostToDevice, pl->stream));
    LB_CUDA(cudaMemcpyAsync(pl->tmpQ.as<u64>() + key_el, evk_a, key_el * 8, cudaMemcpyHostToDevice, pl->stream));
    const uint64_t chunk = std::max<uint64_t>(1, (256ull << 20) / ((in_el + out_el) * 8));
    for (uint64_t off = 0; off < count; off += chunk) {
        const uint64_t m = std::min(chunk, count - off);
        LB_TRY(pl->stage_in.ensure(m * in_el * 8)); LB_TRY(pl->stage_out.ensure(m * out_el * 8));
        LB_CUDA(cudaMemcpyAsync(pl->stage_in.p, digits + off * in_el, m * in_el * 8, cudaMemcpyHostToDevice, pl->stream));
        LB_TRY(luxb200_rns_keyswitch_inner_dev(pl, pl->stage_in.p, pl->tmpQ.p, pl->tmpQ.as<u64>() + key_el, pl->stage_out.p, m, nd, sizeQl));
        LB_CUDA(cudaMemcpyAsync(out + off * out_el, pl->stage_out.p, m * out_el * 8, cudaMemcpyDeviceToHost, pl->stream));
        LB_CUDA(cudaStreamSynchronize(pl->stream));
    }
    return 0;
}

// ---- hybrid key switching of one polynomial: KeySwitchHYBRID::EvalKeySwitchPrecomputeCore (keyswitch-hybrid.cpp:328-412) and
// EvalFastKeySwitchCore (:414-436) as one call over the primitives above.  A digit = `alpha` consecutive limbs of Ql raised to
// Ql u P by ApproxSwitchCRTBasis from its own limbs to all the others (one rns plan per part: source = the part, target = the
// complement), the inner product with the key, ApproxModDown of both results.
int luxb200_hks_plan_create(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, uint32_t alpha,
                            uint32_t sizeQl, int device, luxb200_hks_plan** out) {
    if (!q || !p || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    if (alpha == 0 || sizeQl == 0 || sizeQl > sizeQ || sizeP == 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "need alpha >= 1, 1 <= sizeQl <= sizeQ, sizeP >= 1");
    auto* h = new luxb200_hks_plan();
    h->N = N; h->sizeQ = sizeQ; h->sizeP = sizeP; h->alpha = alpha; h->sizeQl = sizeQl; h->device = device;
    h->nd = std::min((sizeQl + alpha - 1) / alpha, (sizeQ + alpha - 1) / alpha);
    const uint32_t L = sizeQl + sizeP;
    int rc = luxb200_rns_plan_create(N, sizeQ, q, sizeP, p, device, &h->full);
    if (!rc) rc = luxb200_rns_plan_create(N, sizeQl, q, sizeP, p, device, &h->down);
    for (uint32_t part = 0; part < h->nd && !rc; ++part) {
        const uint32_t start = alpha * part, sp = (part == h->nd - 1) ? sizeQl - start : alpha, end = start + sp;
        std::vector<uint64_t> src(q + start, q + end), dst;
        for (uint32_t i = 0; i < L; ++i)
            if (i < start || i >= end) dst.push_back(i < sizeQl ? q[i] : p[i - sizeQl]);
        luxb200_rns_plan* pp = nullptr;
        rc = luxb200_rns_plan_create(N, sp, src.data(), (uint32_t)dst.size(), dst.data(), device, &pp);
        if (!rc && !pp->nttQ) rc = lb_fail(LUXB200_ERR_UNSUPPORTED, "N or the moduli do not admit an NTT");
        h->parts.push_back(pp);
    }
    if (!rc && (!h->full->nttQ || !h->down->nttQ)) rc = lb_fail(LUXB200_ERR_UNSUPPORTED, "N or the moduli do not admit an NTT");
    if (rc) { luxb200_hks_plan_destroy(h); return rc; }
    *out = h;
    return 0;
}
void luxb200_hks_plan_destroy(luxb200_hks_plan* h) {
    if (!h) return;
    for (auto* pp : h->parts) luxb200_rns_plan_destroy(pp);
    luxb200_rns_plan_destroy(h->full);
    luxb200_rns_plan_destroy(h->down);
    delete h;
}
uint32_t luxb200_hks_num_digits(luxb200_hks_plan* h) { return h ? h->nd : 0; }

// the device part for m polynomials: c_dev [m][sizeQl][N], keys [nd][sizeQ + sizeP][N], out_dev [m][2][sizeQl][N]
static int hks_run(luxb200_hks_plan* h, const u64* c_dev, const u64* kb_dev, const u64* ka_dev, u64* out_dev, uint64_t m, uint64_t t) {
    const uint32_t N = h->N, sp_ = h->sizeP, sql = h->sizeQl, nd = h->nd, L = sql + sp_;
    cudaStream_t st = h->full->stream;
    {
        LB_TRY(h->part_in.ensure(m * h->alpha * N * 8)); LB_TRY(h->part_out.ensure(m * L * N * 8));
        LB_TRY(h->digits.ensure(m * nd * L * N * 8)); LB_TRY(h->tilda.ensure(m * 2 * L * N * 8));
        for (uint32_t part = 0; part < nd; ++part) {
            const uint32_t start = h->alpha * part, sp = (part == nd - 1) ? sql - start : h->alpha, end = start + sp;
            auto copy = [&](const u64* in, u32 in_stride, u32 in_off, u64* o, u32 out_stride, u32 out_off, u32 limbs) {
                if (limbs) rns_copy_limbs_kernel<<<nblk(m * limbs * N, 256), 256, 0, st>>>(in, in_stride, in_off, o, out_stride, out_off, limbs, N, m);
            };
            copy(c_dev, sql, start, h->part_in.as<u64>(), sp, 0, sp);
            LB_CUDA(cudaGetLastError());
            LB_CUDA(cudaStreamSynchronize(st));
            LB_TRY(luxb200_approx_mod_up_dev(h->parts[part], h->part_in.p, h->part_out.p, m));   // [own sp | complement L - sp]
            u64* dg = h->digits.as<u64>();
            copy(h->part_out.as<u64>(), L, 0, dg, nd * L, part * L + start, sp);             // own limbs in place (:397-399)
            copy(h->part_out.as<u64>(), L, sp, dg, nd * L, part * L, start);                 // complement before the part (:394-396)
            copy(h->part_out.as<u64>(), L, sp + start, dg, nd * L, part * L + end, L - end); // and after it, P included (:400-402)
            LB_CUDA(cudaGetLastError());
            LB_CUDA(cudaStreamSynchronize(st));
        }
        LB_TRY(luxb200_rns_keyswitch_inner_dev(h->full, h->digits.p, kb_dev, ka_dev, h->tilda.p, m, nd, sql));
        LB_TRY(luxb200_approx_mod_down_bgv_dev(h->down, h->tilda.p, out_dev, 2 * m, t));
    }
    return 0;
}
static uint64_t hks_chunk(const luxb200_hks_plan* h) {
    const uint64_t L = h->sizeQl + h->sizeP;
    const uint64_t per = ((uint64_t)h->sizeQl + 2 * L + (uint64_t)h->nd * L + 2 * L + 2ull * h->sizeQl) * h->N * 8;   // device bytes per polynomial
    return std::max<uint64_t>(1, (8ull << 30) / per);   // up to 8 GB of work buffers: 56 polynomials at N = 2^16 with 24 + 8 limbs
}
// c [count][sizeQl][N] EVALUATION, evk_b / evk_a [nd][sizeQ + sizeP][N], out [count][2][sizeQl][N]; t = 0 CKKS, t > 0 BGV.  Device
// pointers on the plan's device (the key stays resident between calls); returns when the result is complete.
int luxb200_hks_keyswitch_dev(luxb200_hks_plan* h, const void* c_dev, const void* evk_b_dev, const void* evk_a_dev, void* out_dev,
                              uint64_t count, uint64_t t) {
    if (!h || !c_dev || !evk_b_dev || !evk_a_dev || !out_dev) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(h->device));
    const uint64_t chunk = hks_chunk(h), in_el = (uint64_t)h->sizeQl * h->N;
    for (uint64_t off = 0; off < count; off += chunk) {
        const uint64_t m = std::min(chunk, count - off);
        LB_TRY(hks_run(h, reinterpret_cast<const u64*>(c_dev) + off * in_el, reinterpret_cast<const u64*>(evk_b_dev),
                       reinterpret_cast<const u64*>(evk_a_dev), reinterpret_cast<u64*>(out_dev) + off * 2 * in_el, m, t));
    }
    return 0;
}
// the same with host buffers
int luxb200_hks_keyswitch(luxb200_hks_plan* h, const uint64_t* c, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out,
                          uint64_t count, uint64_t t) {
    if (!h || !c || !evk_b || !evk_a || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(h->device));
    const uint32_t N = h->N, sql = h->sizeQl;
    cudaStream_t st = h->full->stream;
    const uint64_t key_el = (uint64_t)h->nd * (h->sizeQ + h->sizeP) * N;
    LB_TRY(h->key.ensure(2 * key_el * 8));
    LB_CUDA(cudaMemcpyAsync(h->key.p, evk_b, key_el * 8, cudaMemcpyHostToDevice, st));
    LB_CUDA(cudaMemcpyAsync(h->key.as<u64>() + key_el, evk_a, key_el * 8, cudaMemcpyHostToDevice, st));
    const uint64_t chunk = hks_chunk(h);
    for (uint64_t off = 0; off < count; off += chunk) {
        const uint64_t m = std::min(chunk, count - off);
        LB_TRY(h->cin.ensure(m * sql * N * 8)); LB_TRY(h->outb.ensure(m * 2 * sql * N * 8));
        LB_CUDA(cudaMemcpyAsync(h->cin.p, c + off * sql * N, m * sql * N * 8, cudaMemcpyHostToDevice, st));
        LB_CUDA(cudaStreamSynchronize(st));
        LB_TRY(hks_run(h, h->cin.as<u64>(), h->key.as<u64>(), h->key.as<u64>() + key_el, h->outb.as<u64>(), m, t));
        LB_CUDA(cudaMemcpyAsync(out + off * 2 * sql * N, h->outb.p, m * 2 * sql * N * 8, cudaMemcpyDeviceToHost, st));
        LB_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

// host-buffer wrappers of the two composites (what = 0 mod up, 1 mod down)
int luxb200_approx_mod_switch(luxb200_rns_plan* pl, const uint64_t* x, uint64_t* out, uint64_t count, int what) {
    return luxb200_approx_mod_switch_bgv(pl, x, out, count, what, 0);
}
int luxb200_approx_mod_switch_bgv(luxb200_rns_plan* pl, const uint64_t* x, uint64_t* out, uint64_t count, int what, uint64_t t) {
    if (!pl || !x || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    if (count == 0) return 0;
    LB_TRY(lb_set_device(pl->device));
    const uint64_t limbs_in = what ? pl->sizeQ + pl->sizeP : pl->sizeQ, limbs_out = what ? pl->sizeQ : pl->sizeQ + pl->sizeP;
    const uint64_t in_el = limbs_in * pl->N, out_el = limbs_out * pl->N;
    const uint64_t chunk = std::max<uint64_t>(1, (256ull << 20) / ((in_el + out_el) * 8));
    for (uint64_t off = 0; off < count; off += chunk) {
        const uint64_t m = std::min(chunk, count - off);
        LB_TRY(pl->stage_in.ensure(m * in_el * 8)); LB_TRY(pl->stage_out.ensure(m * out_el * 8));
        // on the plan's stream and waited for: a synchronous cudaMemcpy from pageable memory may return before its DMA has landed,
        // and the plan's non-blocking streams do not order themselves behind the default stream
        LB_CUDA(cudaMemcpyAsync(pl->stage_in.p, x + off * in_el, m * in_el * 8, cudaMemcpyHostToDevice, pl->stream));
        LB_CUDA(cudaStreamSynchronize(pl->stream));
        if (what) LB_TRY(luxb200_approx_mod_down_bgv_dev(pl, pl->stage_in.p, pl->stage_out.p, m, t));
        else LB_TRY(luxb200_approx_mod_up_dev(pl, pl->stage_in.p, pl->stage_out.p, m));
        LB_CUDA(cudaMemcpyAsync(out + off * out_el, pl->stage_out.p, m * out_el * 8, cudaMemcpyDeviceToHost, pl->stream));
        LB_CUDA(cudaStreamSynchronize(pl->stream));
    }
    return 0;
}

}  // extern "C"
