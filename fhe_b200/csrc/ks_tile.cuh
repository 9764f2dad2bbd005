// ks_tile.cuh — SwitchCTtoqn (lwe-pke.cpp:170-178: ModSwitch Q -> qKS, KeySwitch lwe-pke.cpp:348-372, ModSwitch qKS -> q) for the
// 16-bit switching keys of the N = 1024 parameter sets, organised around the key instead of around the ciphertext.
//
// keyswitch_kernel (aux_kernels.cuh) gives one CTA per ciphertext: it gathers N * digits key rows of npad * 2 bytes each from
// L2 (3.1 MB per STD128 ciphertext, 203 GB per 65 536-gate launch) and is bound by L2 bandwidth; its input stream evicts the
// 99 MB key, so 4 GB come from DRAM for 0.5 GB of algorithmic traffic (profiles/r1_dram_v1.csv).  Here a CTA owns 256
// ciphertexts x 128 key columns with the column sums in registers, and walks the key in coefficient order: the rows of two
// coefficients at a time (2 x baseKS x digits rows x 256 bytes) are staged in shared memory by a three-deep cp.async ring
// and every ciphertext of the CTA adds its `digits` rows per coefficient from there.  The gather moves from L2 (25.8 GB now
// cross it, once per 256 ciphertexts) to shared memory, which is what bounds the kernel (203 GB of conflict-free LDS.64).
//
//   ks_digits_kernel   in [count][N+1] residues mod Qin -> dig [N/2][stride] u32: for coefficients 2p (low half) and 2p+1
//                      (high half) the `digits` base-baseKS digits of RoundqQ(a_i, qKS, Qin), LB bits each
//   ks_tile_kernel     grid (column slices, ciphertext tiles); warp w owns CPW ciphertexts of the tile (CPW w ..), lane l
//                      key columns 4l .. 4l+3 of the slice.  Key entries are < qKS and digits * qKS <= 2^16, so the `digits`
//                      rows of one coefficient are added as packed 16-bit pairs without carry between the halves; the packed
//                      sum s goes into `full` (mod 2^32) and s >> 16 into `hi`, and the even column's sum is full - (hi << 16)
//                      (exact: N * digits * qKS < 2^32).  Slice 0 also gathers the B entries, one lane per (ciphertext, parity).
// Every sum is an exact integer sum reduced mod qKS once, like the reference's ModSubFastEq chain gives the canonical residue:
// the result is bit-identical to keyswitch_kernel and to the reference.
#pragma once
#include "aux_kernels.cuh"

__device__ __forceinline__ void kst_cp16(u32 dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void kst_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_>
__device__ __forceinline__ void kst_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }
__device__ __forceinline__ uint2 kst_lds64(u32 addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// ---- digits, transposed to coefficient-major ------------------------------------------------------------------------------
// block (32, 8): 32 ciphertexts x 64 coefficients per CTA; grid (N / 64, ceil(stride / 32)); ciphertexts >= count give zeros
template <int D, int LB>
__global__ void __launch_bounds__(256) ks_digits_kernel(const u32* __restrict__ in, u32* __restrict__ dig, u32 dig_stride, u32 N,
                                                        u64 qKS, u64 Qin, u32 ms_in, u64 count) {
    __shared__ u32 tile[32][33];   // [ciphertext][coefficient pair], odd pitch: conflict-free both ways
    const u32 i0 = blockIdx.x * 64, c0 = blockIdx.y * 32;
    const u32 lane = threadIdx.x, w = threadIdx.y;
    unsigned short* t16 = reinterpret_cast<unsigned short*>(&tile[0][0]);
#pragma unroll
    for (u32 r = 0; r < 4; ++r) {
        const u32 cl = w * 4 + r;
        const u64 ct = (u64)c0 + cl;
#pragma unroll
        for (u32 h = 0; h < 2; ++h) {
            const u32 il = h * 32 + lane;
            u32 packed = 0;
            if (ct < count) {
                u64 v = in[ct * (N + 1) + i0 + il];
                if (ms_in) v = round_qQ(v, qKS, Qin);
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    packed |= (u32)(v & ((1u << LB) - 1)) << (LB * j);
                    v >>= LB;
                }
            }
            t16[cl * 66 + il] = (unsigned short)packed;
        }
    }
    __syncthreads();
#pragma unroll
    for (u32 r = 0; r < 4; ++r) {
        const u32 p = w * 4 + r;   // coefficient pair of this tile
        dig[(size_t)(i0 / 2 + p) * dig_stride + c0 + lane] = tile[lane][p];
    }
}

// ---- the key walk ------------------------------------------------------------------------------------------------------------
constexpr int KST_CT = 256;      // ciphertexts per CTA
constexpr int KST_COLS = 128;    // key columns per CTA (256 bytes of a row)
constexpr int KST_NST = 3;       // pipeline depth
template <int D, int LB>
constexpr int kst_smem_bytes() { return KST_NST * 2 * (D << LB) * 256 + 8192; }

template <int D, int LB, int CPW>   // CPW ciphertexts per warp: KST_CT / CPW warps per CTA
__global__ void __launch_bounds__(32 * KST_CT / CPW, 1) ks_tile_kernel(const KsParams P, const u32* __restrict__ dig, u32 dig_stride) {
    constexpr int ROWS = D << LB;            // key rows per coefficient
    constexpr int SLAB = ROWS * 256;         // bytes per coefficient in shared memory: [j][a0][256]
    constexpr int STAGE = 2 * SLAB;
    constexpr int CH = 2 * ROWS * 16;        // 16-byte chunks per stage
    constexpr int NT = 32 * KST_CT / CPW;    // threads
    constexpr int CPT = (CH + NT - 1) / NT;  // chunks per thread
    extern __shared__ unsigned char kst_raw[];
    // stage bases are multiples of 8 KB, so that (digit << 8) can be OR-ed into a lane's base address
    const u32 sm0 = ((u32)__cvta_generic_to_shared(kst_raw) + 8191u) & ~8191u;
    static_assert(SLAB % 8192 == 0, "slab must keep the 8 KB alignment");
    const u32 slice = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const u64 ct0 = (u64)blockIdx.y * KST_CT + warp * CPW;
    const u32 rowbytes = P.npad * 2;
    const u32 NS = P.N / 2;

    // this thread's chunks of a stage: global row r = ii * ROWS + a0 * D + j (key order), shared row ii * ROWS + j * 2^LB + a0
    u32 dsto[CPT];
    u32 srco[CPT];
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const u32 q = tid + NT * k, r = q >> 4, c = q & 15;
        const u32 ii = r / ROWS, rr = r % ROWS, a0 = rr / D, j = rr % D;
        const bool ok = q < CH && slice * 256 + c * 16 < rowbytes;
        dsto[k] = ok ? (ii * ROWS + (j << LB) + a0) * 256 + c * 16 : 0xffffffffu;
        srco[k] = r * rowbytes + slice * 256 + c * 16;
    }
    const unsigned char* kA = reinterpret_cast<const unsigned char*>(P.kA);
    auto issue = [&](u32 st) {
        const u32 dst = sm0 + (st % KST_NST) * STAGE;
        const unsigned char* src = kA + (size_t)st * 2 * ROWS * rowbytes;
#pragma unroll
        for (int k = 0; k < CPT; ++k)
            if (dsto[k] != 0xffffffffu) kst_cp16(dst + dsto[k], src + srco[k]);
    };

    u32 full[CPW][2], hi[CPW][2];
#pragma unroll
    for (int t = 0; t < CPW; ++t) full[t][0] = full[t][1] = hi[t][0] = hi[t][1] = 0;
    u32 bs = 0, bpend[D];
#pragma unroll
    for (int j = 0; j < D; ++j) bpend[j] = 0;
    const u32* dptr = dig + ct0 + (lane & (CPW - 1));
    const u32 bpar = (lane / CPW) & 1;   // slice 0: lane groups of CPW alternate between the even and the odd coefficient

    issue(0); kst_commit();
    if (NS > 1) issue(1);
    kst_commit();
    u32 dnext = dptr[0];
    for (u32 st = 0; st < NS; ++st) {
        kst_wait<KST_NST - 2>();
        __syncthreads();
        if (st + KST_NST - 1 < NS) issue(st + KST_NST - 1);
        kst_commit();
        const u32 dreg = dnext;
        if (st + 1 < NS) dnext = dptr[(size_t)(st + 1) * dig_stride];
        if (slice == 0 && lane < 2 * CPW) {   // the gathered entries are added a stage later: nothing waits on the loads
            const u32 dd = dreg >> (16 * bpar);
            const u32 i = 2 * st + bpar;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const u32 a0 = (dd >> (LB * j)) & ((1u << LB) - 1);
                bs += bpend[j];
                bpend[j] = P.kB32[((size_t)i * (1u << LB) + a0) * D + j];
            }
        }
        const u32 lbase = sm0 + (st % KST_NST) * STAGE + lane * 8;
#pragma unroll
        for (int t = 0; t < CPW; ++t) {
            const u32 d = __shfl_sync(0xffffffffu, dreg, t);
#pragma unroll
            for (int ii = 0; ii < 2; ++ii) {
                u32 s0 = 0, s1 = 0;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    constexpr u32 MASK = ((1u << LB) - 1) << 8;
                    const int sh = 16 * ii + LB * j - 8;
                    const u32 x = sh >= 0 ? d >> (sh >= 0 ? sh : 0) : d << (sh < 0 ? -sh : 0);
                    const uint2 v = kst_lds64(((x & MASK) | (lbase + ii * SLAB)) + j * (256 << LB));
                    s0 += v.x; s1 += v.y;
                }
                full[t][0] += s0; hi[t][0] += s0 >> 16;
                full[t][1] += s1; hi[t][1] += s1 >> 16;
            }
        }
    }
    kst_wait<0>();

    const u32 qKS = (u32)P.qKS;
    const u32 k0 = slice * KST_COLS + lane * 4;
#pragma unroll
    for (int t = 0; t < CPW; ++t) {
        const u64 ct = ct0 + t;
        u32 o[4];
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            const u32 sums[2] = {full[t][w] - (hi[t][w] << 16), hi[t][w]};
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const u32 s = sums[e] % qKS;
                u64 v = s ? qKS - s : 0;   // a = 0 - sum
                if (P.ms_out) v = round_qQ(v, P.fmod, P.qKS);
                o[2 * w + e] = (u32)v;
            }
        }
        u32* oa = reinterpret_cast<u32*>(P.out_a) + ct * P.out_stride;
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (ct < P.count && k0 + e < P.n) oa[k0 + e] = o[e];
    }
    if (slice == 0) {
#pragma unroll
        for (int j = 0; j < D; ++j) bs += bpend[j];
        bs += __shfl_xor_sync(0xffffffffu, bs, CPW);
        const u64 ct = ct0 + lane;
        if (lane < CPW && ct < P.count) {
            u64 b = reinterpret_cast<const u32*>(P.in)[ct * (P.N + 1) + P.N];
            if (P.ms_in) b = round_qQ(b, P.qKS, P.Qin);
            b = (b + P.qKS - bs % qKS) % P.qKS;
            if (P.ms_out) b = round_qQ(b, P.fmod, P.qKS);
            reinterpret_cast<u32*>(P.out_b)[ct] = (u32)b;
        }
    }
}
