// api_common.h — host-side context objects shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/luxb200.h"
#include "br_common.cuh"
#include "host_math.h"

extern thread_local std::string g_last_error;
int lb_fail(int code, const char* fmt, ...);

#define LB_CUDA(expr)                                                                                  \
    do {                                                                                               \
        cudaError_t e_ = (expr);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            return lb_fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? LUXB200_ERR_NO_DEVICE \
                                                                                        : LUXB200_ERR_CUDA,    \
                           "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)
#define LB_TRY(expr)            \
    do {                        \
        int r_ = (expr);        \
        if (r_ != 0) return r_; \
    } while (0)

// grow-only device buffer; owns its allocation (freed on destruction, so early error returns do not leak)
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool borrowed = false;   // alias of another buffer's allocation (luxb200_share_ksk): never freed here
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), cap(o.cap), borrowed(o.borrowed) { o.p = nullptr; o.cap = 0; o.borrowed = false; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; cap = o.cap; borrowed = o.borrowed; o.p = nullptr; o.cap = 0; o.borrowed = false; }
        return *this;
    }
    void borrow(const DevBuf& o) { release(); p = o.p; cap = o.cap; borrowed = true; }
    ~DevBuf() { release(); }
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        release();
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) return lb_fail(LUXB200_ERR_ALLOC, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
        cap = bytes;
        return 0;
    }
    void release() {
        if (p && !borrowed) cudaFree(p);
        p = nullptr;
        cap = 0;
        borrowed = false;
    }
    template <typename T>
    T* as() { return reinterpret_cast<T*>(p); }
};

// u32 NTT tables of one modulus (device) + the uniform constants
struct Ntt32Tables {
    uint32_t N = 0, logN = 0, logH = 0;
    uint64_t Q = 0, psi = 0;
    NttConsts C;
    DevBuf twB, twiB, twfull, twifull;
    int build(uint32_t N, uint64_t Q, uint64_t psi);
    int build_half(uint32_t N, uint64_t Q, const std::vector<uint64_t>& tw2, const std::vector<uint64_t>& twi2, uint32_t h, uint64_t ninv2);
    int build_from(uint32_t N, uint64_t Q, const std::vector<uint64_t>& tw, const std::vector<uint64_t>& twi, uint64_t ninv);
    void release() { twB.release(); twiB.release(); twfull.release(); twifull.release(); }
};

struct luxb200_ctx {
    // One context = one device, one stream, one set of work buffers: the entry points serialise on this mutex, so several
    // host threads (the reference's std::async batch futures, batch.cpp:320-349) may share a context safely; calls on
    // DIFFERENT contexts run concurrently.
    std::recursive_mutex mu;
    luxb200_params p;
    int device = 0;
    uint32_t digitsG = 0, digitsKS = 0, digitsR = 0, rows = 0, logN = 0;
    uint32_t npad = 0;       // padded KSK row length
    uint32_t ks_elem = 0;    // bytes per KSK element (2 or 4)
    uint32_t red_s = 0, red_mu = 0, red_mu_hi = 0, red_r = 0, red_rs = 0;
    // br_quad.cuh (N = 2048 as four warps per ciphertext): `ntt` holds the twiddle set of half 0, `ntt1` of half 1
    bool quad = false;
    Ntt32Tables ntt1;
    uint32_t xw = 0, xws = 0, xwi = 0, xwis = 0;   // cross-stage twiddles table[1], inverse table[1] (with Shoup companions)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    Ntt32Tables ntt;
    // "wide" contexts (br_wide.cuh): N = 2048, 64-bit Q, or a digit count the register-resident kernels are not built for
    struct Wide {
        bool on = false;
        uint32_t wsize = 4;                    // bytes per residue on the device: 4 (Q < 2^30) or 8
        DevBuf tw, twi, pw, automap;           // (w, Shoup) tables [N] x 2, psi powers [2N], natural-order automorphism maps
        uint64_t ninv = 0, ninv_s = 0, mu = 0, r = 0, rs = 0;
        size_t smem = 0;                       // dynamic shared memory of brw_kernel
    } wide;
    DevBuf bsk, mono, kskA, kskB, automap, loggen;
    bool have_bsk = false, have_ksk = false;
    // work buffers
    DevBuf in0, in1, in2, in3, wa, wb, lut, ext, oa, ob, o64a, o64b, sched, tmpa, tmpb, tmpc, tmpd;
    DevBuf cx1a, cx1b, cx2a, cx2b;   // CMUX intermediates (two NAND results), kept so the call does no cudaMalloc/cudaFree
    DevBuf xin;                      // extended (N, Q) inputs packed [count][N+1] before SwitchCTtoqn
    DevBuf ksdig;                    // ks_tile.cuh: key-switch digits, coefficient-major [N/2][stride] u32
    DevBuf cpa, cpb, cdesc, clsel;   // gate circuits (circuit.cu): wire pool (a, b), gate descriptors, per-item table offsets
    uint64_t launches = 0;
    float t_prep = 0, t_br = 0, t_ks = 0, t_total = 0;
    uint32_t last_launches = 0;
    uint64_t circ_stats[3] = {0, 0, 0};   // levels, bootstrapped gates, launches of the last circuit call
    // host-buffer pipeline (luxb200_evalbingate_batch with large batches): copy-in / copy-out streams, two input landing
    // sets and two output staging sets, so the PCIe copies of chunk k+1 / k-1 run under the kernels of chunk k
    struct HostPipe {
        cudaStream_t s_in = nullptr, s_out = nullptr;
        cudaEvent_t e_in[2] = {}, e_free[2] = {}, e_stage[2] = {}, e_out[2] = {};
        DevBuf in[2][4], st[2][2];
        std::vector<cudaEvent_t> tev;   // 4 timing events per chunk
        bool ready = false;
    } pipe;
};

int lb_set_device(int device);
int lb_launch_quad(luxb200_ctx* c, const BrParams& P, int sched_only);   // br11.cu
int lb_build_wide_tables(luxb200_ctx* c);
#define LB_LOCK(c) std::lock_guard<std::recursive_mutex> lb_guard_((c)->mu)

// ---- entry points of api.cu used by circuit.cu (a separate translation unit, same library) -------------------
// test-vector table of a 2-input gate for plaintext modulus p (binfhe-base-scheme.cpp:535-567)
void lb_gate_lut(const luxb200_ctx* c, int gate, uint64_t p, std::vector<uint64_t>& lut);
int lb_upload_lut(luxb200_ctx* c, const std::vector<uint64_t>& lut);
// blind rotation + extraction + SwitchCTtoqn on prepared device inputs; lutsel: optional per-item offsets into c->lut
int lb_boot_dev(luxb200_ctx* c, const uint32_t* d_a, const uint32_t* d_b, uint64_t ctmod, uint64_t b_add, uint64_t fmod,
                uint32_t* d_oa, uint32_t* d_ob, uint64_t count, const uint32_t* lutsel);
int lb_download_rows(luxb200_ctx* c, const uint32_t* d_in, uint32_t in_stride, uint32_t in_off, uint32_t len, uint64_t count,
                     DevBuf& stage, uint64_t* host);
uint64_t lb_wave(luxb200_ctx* c);   // ciphertexts one blind-rotation wave holds (SMs x ciphertexts per CTA)
