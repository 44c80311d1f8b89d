/*
 * oracle/ref_harness.c -- TEST INFRASTRUCTURE ONLY (see oracle/edt_oracle.c header).
 *
 * Builds the UNMODIFIED reference hot path into oracle/_ref/libchaqref.so by #including the reference's
 * own translation units from where they lie in the reference tree (REF_ROOT, default /root/reference).
 * No reference source is copied into this repository; oracle/_ref/ is git-ignored.
 *
 * What is exposed (all on raw buffers, no file I/O):
 *   ref_sdf_u8            = sdfgen.c:264-300 verbatim call sequence (a1, a2+a6 twice in nested OpenMP
 *                           sections, a8, a9), using the reference's own static functions.
 *   ref_dist_transform_2d = df.h:6.
 *   ref_d2_field          = the pre-sqrt squared field: dist_transform_axis(.., false) twice (df.c:130 and
 *                           the same call in place of df.c:133), to compare squared distances directly.
 *   ref_load_ya           = stbi_load(..., req_comp = 2) as sdfgen.c:250-258 does, for making fixtures.
 *
 * Limits of the stock code that callers must respect: square images only
 * (df.c:47/:92 stride bug), exact only for side <= 4098.
 */
#include <omp.h>

#define main chaq_ref_main
#include "openmp/sdfgen.c" /* resolved through -I$(REF_ROOT) */
#undef main
#include "openmp/df.c"

int ref_sdf_u8(const unsigned char* img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
               size_t spread, int asymmetric, unsigned char* out) {
    omp_set_nested(1); /* sdfgen.c:123 */
    bool* img_bool = malloc(w * h * sizeof(bool));
    float* fin = malloc(w * h * sizeof(float));
    float* fout = malloc(w * h * sizeof(float));
    if (!img_bool || !fin || !fout) return -1;
    transform_img_to_bool(img, img_bool, w, h, stride, offset, test_above != 0);
#pragma omp parallel sections num_threads(2)
    {
#pragma omp section
        {
            transform_bool_to_float(img_bool, fin, w, h, true);
            dist_transform_2d(fin, w, h);
        }
#pragma omp section
        {
            transform_bool_to_float(img_bool, fout, w, h, false);
            dist_transform_2d(fout, w, h);
        }
    }
    transform_float_sub(fout, fin, w, h);
    transform_float_to_byte(fout, out, w, h, spread, asymmetric != 0);
    free(img_bool); free(fin); free(fout);
    return 0;
}

/* Same sequence, keeping the float fields. Any output may be NULL. */
int ref_sdf_fields(const unsigned char* img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                   size_t spread, int asymmetric, unsigned char* out, float* sdf, float* d_in, float* d_out) {
    omp_set_nested(1);
    bool* img_bool = malloc(w * h * sizeof(bool));
    float* fin = malloc(w * h * sizeof(float));
    float* fout = malloc(w * h * sizeof(float));
    if (!img_bool || !fin || !fout) return -1;
    transform_img_to_bool(img, img_bool, w, h, stride, offset, test_above != 0);
    transform_bool_to_float(img_bool, fin, w, h, true);
    dist_transform_2d(fin, w, h);
    transform_bool_to_float(img_bool, fout, w, h, false);
    dist_transform_2d(fout, w, h);
    if (d_in) memcpy(d_in, fin, w * h * sizeof(float));
    if (d_out) memcpy(d_out, fout, w * h * sizeof(float));
    transform_float_sub(fout, fin, w, h);
    if (sdf) memcpy(sdf, fout, w * h * sizeof(float));
    if (out) transform_float_to_byte(fout, out, w, h, spread, asymmetric != 0);
    free(img_bool); free(fin); free(fout);
    return 0;
}

void ref_dist_transform_2d(float* img, size_t w, size_t h) { dist_transform_2d(img, w, h); }

/* squared field (no sqrt). img in/out, square images only. */
void ref_d2_field(float* img, size_t w, size_t h) {
    float* tp = malloc(w * h * sizeof(float));
    dist_transform_axis(img, w, h, tp, false);
    dist_transform_axis(tp, h, w, img, false);
    free(tp);
}

/* Decode exactly as the CLI does: 2 channels (Y, A). Returns malloc'd buffer (free with ref_free). */
unsigned char* ref_load_ya(const char* path, int* w, int* h, int* n) { return stbi_load(path, w, h, n, 2); }
void ref_free(void* p) { free(p); }

int ref_omp_max_threads(void) { return omp_get_max_threads(); }
