// aux_kernels.cuh — key packing, LWE pre-processing, ModSwitch + KeySwitch + ModSwitch.
#pragma once
#include "modarith.cuh"

// ---- key packing -------------------------------------------------------------------------------------------
// canonical u64 polynomials (EVALUATION order p) -> u32, sigma order: p = E*l + r  ->  (r/4)*128 + 4*l + (r%4)
// quad = 1 (br_quad.cuh): a polynomial is two halves of N/2 evaluation points, each in the sigma order of its own indices
__global__ void pack_polys_sigma_kernel(const u64* __restrict__ in, u32* __restrict__ out, u32 logN, u64 total, u32 quad = 0) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const u32 N = 1u << logN, S = N >> quad, E = S >> 5;
    const u64 poly = idx >> logN;
    const u32 p = (u32)(idx & (N - 1)), ps = p & (S - 1);
    const u32 l = ps / E, r = ps % E;
    out[poly * N + (p - ps) + (r >> 2) * 128 + 4 * l + (r & 3)] = (u32)in[idx];
}
__global__ void unpack_polys_sigma_kernel(const u32* __restrict__ in, u64* __restrict__ out, u32 logN, u64 total, u32 quad = 0) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const u32 N = 1u << logN, S = N >> quad, E = S >> 5;
    const u64 poly = idx >> logN;
    const u32 p = (u32)(idx & (N - 1)), ps = p & (S - 1);
    const u32 l = ps / E, r = ps % E;
    out[idx] = in[poly * N + (p - ps) + (r >> 2) * 128 + 4 * l + (r & 3)];
}
// KSK A: u64 [rows][n] -> KT [rows][npad] (zero padded)
template <typename KT>
__global__ void pack_ksk_kernel(const u64* __restrict__ in, KT* __restrict__ out, u32 n, u32 npad, u64 rows) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * npad) return;
    const u64 row = idx / npad;
    const u32 k = (u32)(idx % npad);
    out[idx] = k < n ? (KT)in[row * n + k] : (KT)0;
}
template <typename KT>
__global__ void unpack_ksk_kernel(const KT* __restrict__ in, u64* __restrict__ out, u32 n, u32 npad, u64 rows) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * n) return;
    const u64 row = idx / n;
    const u32 k = (u32)(idx % n);
    out[idx] = (u64)in[row * npad + k];
}
__global__ void narrow_u64_u32_kernel(const u64* __restrict__ in, u32* __restrict__ out, u64 total) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < total) out[idx] = (u32)in[idx];
}
__global__ void copy_u64_kernel(const u64* __restrict__ in, u64* __restrict__ out, u64 total) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < total) out[idx] = in[idx];
}
__global__ void widen_u32_u64_kernel(const u32* __restrict__ in, u64* __restrict__ out, u64 total) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < total) out[idx] = in[idx];
}

// ---- LWE linear pre-processing (EvalAddEq / EvalAddConstEq, lwe-pke.cpp:225-251) ----------------------------
// out = mul * (sum_{j<k} in[ct][j][:]) + add  (mod q), for the (n+1)-vector (a, b).
// IT = u64 (host-format inputs) or u32 (device-resident inputs).  in_a: [count][k][n], in_b: [count][k]
template <typename IT>
__global__ void lwe_combine_kernel(const IT* __restrict__ in_a, const IT* __restrict__ in_b, const IT* __restrict__ in_a2,
                                   const IT* __restrict__ in_b2, u32* __restrict__ out_a, u32* __restrict__ out_b, u32 n,
                                   u32 k, u32 q, u32 mul, u32 addb, u64 count) {
    // two calling shapes: (in_a2 == nullptr) k-way sum from one array; else 2 inputs from two arrays
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 per = (u64)n + 1;
    if (idx >= count * per) return;
    const u64 ct = idx / per;
    const u32 t = (u32)(idx % per);
    u64 s = 0;
    if (in_a2) {
        s = (t < n) ? ((u64)in_a[ct * n + t] % q + (u64)in_a2[ct * n + t] % q) : ((u64)in_b[ct] % q + (u64)in_b2[ct] % q);
    } else {
        for (u32 j = 0; j < k; ++j) s += (t < n) ? (u64)in_a[(ct * k + j) * n + t] % q : (u64)in_b[ct * k + j] % q;
    }
    s = (s % q) * mul % q;
    if (t < n) out_a[ct * n + t] = (u32)s;
    else out_b[ct] = (u32)((s + addb) % q);
}

// EvalNOT (binfhe-base-scheme.cpp:223-237): a -> -a, b -> q/4 - b
__global__ void lwe_not_kernel(const u32* __restrict__ a, const u32* __restrict__ b, u32* __restrict__ oa,
                               u32* __restrict__ ob, u32 n, u32 q, u64 count) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 per = (u64)n + 1;
    if (idx >= count * per) return;
    const u64 ct = idx / per;
    const u32 t = (u32)(idx % per);
    if (t < n) {
        u32 v = a[ct * n + t];
        oa[ct * n + t] = v ? q - v : 0;
    } else {
        u32 v = b[ct], q4 = q >> 2;
        ob[ct] = q4 >= v ? q4 - v : q4 + q - v;
    }
}

// elementwise RoundqQ
__global__ void modswitch_kernel(const u64* __restrict__ in, u64* __restrict__ out, u64 qto, u64 qfrom, u64 len) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < len) out[idx] = round_qQ(in[idx], qto, qfrom);
}

// ---- SwitchCTtoqn (lwe-pke.cpp:170-178): ModSwitch(Q->qKS) + KeySwitch (lwe-pke.cpp:348-372) + ModSwitch(qKS->fmod)
struct KsParams {
    const void* in;     // IN [count][N+1]  (a then b), mod Qin
    const void* kA;     // KT [N*baseKS*digits][npad]
    const u32* kB32;    // [N*baseKS*digits]  (qKS <= 2^32)
    const u64* kB64;    //                    (qKS >  2^32)
    void* out_a;        // [count][out_stride] first n entries; u32, or u64 when out64
    void* out_b;        // [count]
    u64 Qin, qKS, fmod;
    u32 N, n, npad, baseKS, digits, out_stride;
    u32 ms_in, ms_out;  // apply RoundqQ on input (Qin->qKS) / output (qKS->fmod)
    u32 out64, pad;
    u64 count;
};

// KT: key element (u16 / u32 / u64 by the size of qKS), ACC: column accumulator (N * digits * qKS must fit),
// IN: residue type of the extracted ciphertexts (u32, or u64 for the 64-bit blind rotation)
template <typename KT, typename ACC, typename IN>
__global__ void __launch_bounds__(128) keyswitch_kernel(const KsParams P) {
    extern __shared__ u32 rowidx[];  // [N*digits]
    __shared__ u64 bred[128];
    const u64 ct = blockIdx.x;
    const IN* in = reinterpret_cast<const IN*>(P.in) + ct * (P.N + 1);
    const u32 R = P.N * P.digits;
    for (u32 i = threadIdx.x; i < P.N; i += blockDim.x) {
        u64 v = in[i];
        if (P.ms_in) v = round_qQ(v, P.qKS, P.Qin);
        for (u32 j = 0; j < P.digits; ++j) {
            const u32 a0 = (u32)(v % P.baseKS);
            v /= P.baseKS;
            rowidx[i * P.digits + j] = (i * P.baseKS + a0) * P.digits + j;
        }
    }
    __syncthreads();
    constexpr int VEC = 8 / sizeof(KT);
    const KT* kA = reinterpret_cast<const KT*>(P.kA);
    // b: each thread sums a strided subset of the B entries (modular: the entries may be up to 2^62)
    u64 bs = 0;
    for (u32 r = threadIdx.x; r < R; r += blockDim.x) {
        bs += P.kB64 ? P.kB64[rowidx[r]] : (u64)P.kB32[rowidx[r]];
        if (bs >= P.qKS) bs %= P.qKS;
    }
    bred[threadIdx.x] = bs;
    for (u32 k0 = threadIdx.x * VEC; k0 < P.npad; k0 += blockDim.x * VEC) {
        ACC acc[VEC];
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[e] = 0;
        u32 r = 0;
        for (; r + 8 <= R; r += 8) {
            uint2 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                v[u] = __ldg(reinterpret_cast<const uint2*>(kA + (size_t)rowidx[r + u] * P.npad + k0));
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (VEC == 4) {
                    acc[0] += v[u].x & 0xffffu;
                    acc[1 % VEC] += v[u].x >> 16;
                    acc[2 % VEC] += v[u].y & 0xffffu;
                    acc[3 % VEC] += v[u].y >> 16;
                } else if (VEC == 2) {
                    acc[0] += v[u].x;
                    acc[1 % VEC] += v[u].y;
                } else {
                    acc[0] += (ACC)((u64)v[u].x | ((u64)v[u].y << 32));
                }
            }
        }
        for (; r < R; ++r) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(kA + (size_t)rowidx[r] * P.npad + k0));
            if (VEC == 4) {
                acc[0] += v.x & 0xffffu;
                acc[1 % VEC] += v.x >> 16;
                acc[2 % VEC] += v.y & 0xffffu;
                acc[3 % VEC] += v.y >> 16;
            } else if (VEC == 2) {
                acc[0] += v.x;
                acc[1 % VEC] += v.y;
            } else {
                acc[0] += (ACC)((u64)v.x | ((u64)v.y << 32));
            }
        }
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            const u32 k = k0 + e;
            if (k < P.n) {
                u64 s = (u64)acc[e] % P.qKS;
                u64 v = s ? P.qKS - s : 0;  // a = 0 - sum
                if (P.ms_out) v = round_qQ(v, P.fmod, P.qKS);
                if (P.out64) reinterpret_cast<u64*>(P.out_a)[ct * P.out_stride + k] = v;
                else reinterpret_cast<u32*>(P.out_a)[ct * P.out_stride + k] = (u32)v;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        u64 s = 0;
        for (int i = 0; i < 128; ++i) { s += bred[i]; if (s >= P.qKS) s %= P.qKS; }
        s %= P.qKS;
        u64 b = in[P.N];
        if (P.ms_in) b = round_qQ(b, P.qKS, P.Qin);
        b = (b + P.qKS - s) % P.qKS;
        if (P.ms_out) b = round_qQ(b, P.fmod, P.qKS);
        if (P.out64) reinterpret_cast<u64*>(P.out_b)[ct] = b;
        else reinterpret_cast<u32*>(P.out_b)[ct] = (u32)b;
    }
}

// ---- integer-pipe probe -------------------------------------------------------------------------------------
// imad: 8 independent dependent-chains of IMAD per thread; mixed: the Shoup butterfly instruction mix.
__global__ void intpipe_imad_kernel(u32* out, u32 iters, u32 seed) {
    u32 a0 = seed + threadIdx.x, a1 = a0 * 3 + 1, a2 = a0 * 5 + 2, a3 = a0 * 7 + 3, a4 = a0 * 9 + 4, a5 = a0 * 11 + 5,
        a6 = a0 * 13 + 6, a7 = a0 * 15 + 7;
    const u32 m = seed | 1u, c = seed ^ 0x9e3779b9u;
    for (u32 i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            a0 = a0 * m + c; a1 = a1 * m + c; a2 = a2 * m + c; a3 = a3 * m + c;
            a4 = a4 * m + c; a5 = a5 * m + c; a6 = a6 * m + c; a7 = a7 * m + c;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
}
__global__ void intpipe_bfly_kernel(u32* out, u32 iters, u32 seed) {
    u32 x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) x[u] = seed * (u + 1) + threadIdx.x;
    const u32 Q = 134215681u, Q2 = 2 * Q, w = (seed % Q) | 1u, ws = (u32)(((u64)w << 32) / Q);
    for (u32 i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; u += 2) {
            u32 t = shoup_mul(x[u + 1], w, ws, Q);
            x[u + 1] = csub(x[u] - t + Q2, Q2);
            x[u] = csub(x[u] + t, Q2);
        }
    }
    u32 r = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) r ^= x[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
