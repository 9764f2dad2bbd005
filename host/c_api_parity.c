/* c_api_parity.c — plain C, through the FFI only: the reference's lux_fhe_* calls create the context, keys and
 * ciphertexts; lux_fhe_*_batch (B200) is compared with the reference's single-ciphertext calls (CPU) byte for byte via
 * lux_fhe_ciphertext_marshal.  TEST INFRASTRUCTURE.  usage: c_api_parity [toy|std128] [count] */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "lux/fhe/c_api_batch.h"

static int same(const LuxFheCiphertext* x, const LuxFheCiphertext* y) {
    uint8_t *dx = NULL, *dy = NULL;
    size_t lx = 0, ly = 0;
    if (lux_fhe_ciphertext_marshal(x, &dx, &lx) != LUX_FHE_OK || lux_fhe_ciphertext_marshal(y, &dy, &ly) != LUX_FHE_OK) return 0;
    int ok = lx == ly && memcmp(dx, dy, lx) == 0;
    lux_fhe_bytes_free(dx);
    lux_fhe_bytes_free(dy);
    return ok;
}

typedef LuxFheError (*gate1_fn)(LuxFheContext*, const LuxFheBootstrapKey*, const LuxFheCiphertext*, const LuxFheCiphertext*, LuxFheCiphertext**);
typedef LuxFheError (*gateb_fn)(LuxFheContext*, const LuxFheBootstrapKey*, const LuxFheCiphertext* const*, const LuxFheCiphertext* const*, size_t, LuxFheCiphertext**);

int main(int argc, char** argv) {
    const int std128 = argc > 1 && strcmp(argv[1], "std128") == 0;
    const size_t n = argc > 2 ? (size_t)atoi(argv[2]) : 8;
    LuxFheContext* ctx = NULL; LuxFheSecretKey* sk = NULL; LuxFheBootstrapKey* bsk = NULL;
    /* LUX_FHE_PARAMS_STD128 maps to STD128_LMKCDEY for every method (c_api.cpp:46-58) */
    if (lux_fhe_context_new(&ctx, std128 ? LUX_FHE_PARAMS_STD128 : LUX_FHE_PARAMS_TOY,
                            std128 ? LUX_FHE_METHOD_LMKCDEY : LUX_FHE_METHOD_GINX) != LUX_FHE_OK) return 2;
    if (lux_fhe_keygen_secret(ctx, &sk) != LUX_FHE_OK || lux_fhe_keygen_bootstrap(ctx, sk, &bsk) != LUX_FHE_OK) return 2;

    LuxFheCiphertext **a = calloc(n, sizeof(*a)), **b = calloc(n, sizeof(*b)), **s = calloc(n, sizeof(*s)), **out = calloc(n, sizeof(*out));
    int fails = 0;
    for (size_t i = 0; i < n; ++i) {
        lux_fhe_encrypt(ctx, sk, (i >> 0) & 1, &a[i]);
        lux_fhe_encrypt(ctx, sk, (i >> 1) & 1, &b[i]);
        lux_fhe_encrypt(ctx, sk, (i >> 2) & 1, &s[i]);
    }
    /* before lux_fhe_gpu_enable the batch calls must refuse */
    if (lux_fhe_and_batch(ctx, bsk, (const LuxFheCiphertext* const*)a, (const LuxFheCiphertext* const*)b, n, out) != LUX_FHE_ERR_NOT_INIT) {
        printf("and_batch did not refuse before gpu_enable\n"); ++fails;
    }
    LuxFheError e = lux_fhe_gpu_enable(ctx, bsk, 0);
    if (e != LUX_FHE_OK) { printf("lux_fhe_gpu_enable: %s (no CUDA device: no CPU fallback)\n", lux_fhe_strerror(e)); return 3; }

    const char* names[6] = {"and", "or", "xor", "nand", "nor", "xnor"};
    gate1_fn g1[6] = {lux_fhe_and, lux_fhe_or, lux_fhe_xor, lux_fhe_nand, lux_fhe_nor, lux_fhe_xnor};
    gateb_fn gb[6] = {lux_fhe_and_batch, lux_fhe_or_batch, lux_fhe_xor_batch, lux_fhe_nand_batch, lux_fhe_nor_batch, lux_fhe_xnor_batch};
    for (int g = 0; g < 6; ++g) {
        size_t bad = 0;
        if (gb[g](ctx, bsk, (const LuxFheCiphertext* const*)a, (const LuxFheCiphertext* const*)b, n, out) != LUX_FHE_OK) { printf("%s_batch failed\n", names[g]); return 2; }
        for (size_t i = 0; i < n; ++i) {
            LuxFheCiphertext* r = NULL;
            g1[g](ctx, bsk, a[i], b[i], &r);
            bad += !same(r, out[i]);
            lux_fhe_ciphertext_free(r);
            lux_fhe_ciphertext_free(out[i]);
        }
        printf("lux_fhe_%s_batch: %zu compared, %zu mismatches\n", names[g], n, bad);
        fails += bad != 0;
    }
    {
        size_t bad = 0;
        if (lux_fhe_mux_batch(ctx, bsk, (const LuxFheCiphertext* const*)s, (const LuxFheCiphertext* const*)a, (const LuxFheCiphertext* const*)b, n, out) != LUX_FHE_OK) return 2;
        for (size_t i = 0; i < n; ++i) {
            LuxFheCiphertext* r = NULL;
            lux_fhe_mux(ctx, bsk, s[i], a[i], b[i], &r);
            bad += !same(r, out[i]);
            lux_fhe_ciphertext_free(r); lux_fhe_ciphertext_free(out[i]);
        }
        printf("lux_fhe_mux_batch: %zu compared, %zu mismatches\n", n, bad);
        fails += bad != 0;
    }
    {
        size_t bad = 0;
        if (lux_fhe_bootstrap_batch(ctx, bsk, (const LuxFheCiphertext* const*)a, n, out) != LUX_FHE_OK) return 2;
        for (size_t i = 0; i < n; ++i) {
            LuxFheCiphertext* r = NULL;
            bool bit = 0;
            lux_fhe_bootstrap(ctx, bsk, a[i], &r);
            bad += !same(r, out[i]);
            lux_fhe_decrypt(ctx, sk, out[i], &bit);
            bad += bit != (bool)(i & 1);
            lux_fhe_ciphertext_free(r); lux_fhe_ciphertext_free(out[i]);
        }
        printf("lux_fhe_bootstrap_batch: %zu compared, %zu mismatches\n", n, bad);
        fails += bad != 0;
    }
    printf("{\"c_api_parity\": \"%s\", \"failed_checks\": %d}\n", std128 ? "std128" : "toy", fails);
    return fails ? 1 : 0;
}
