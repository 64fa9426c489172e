/* sift3d_oracle.h -- CPU restatement of the reference's detect + describe path.
 * TEST INFRASTRUCTURE ONLY (see sift3d_oracle.c). */
#ifndef SIFT3D_ORACLE_H
#define SIFT3D_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_EXPORT __attribute__((visibility("default")))
#define ORC_DESC_NUMEL 768

typedef struct orc_detector orc_detector;

/* Same layouts as oracle/ref_stage_harness.c's refstage_cand / refstage_key. */
typedef struct {
    int32_t o, s;
    double x, y, z, sd;
    float strength;
    int32_t pad;
} orc_cand;

typedef struct {
    int32_t o, s;
    double x, y, z, sd;
    float strength;
    float R[9];
    int32_t pad;
} orc_key;

ORC_EXPORT orc_detector *orc_make(void);
ORC_EXPORT void orc_free(orc_detector *);
ORC_EXPORT int orc_set_params(orc_detector *, double peak_thresh, double corner_thresh,
                              int num_kp_levels, double sigma_n, double sigma0);
ORC_EXPORT int orc_detect(orc_detector *, const float *vol, int nx, int ny, int nz, double ux,
                          double uy, double uz);
ORC_EXPORT int orc_num_octaves(const orc_detector *);
ORC_EXPORT int orc_num_levels(const orc_detector *);
ORC_EXPORT const float *orc_level(const orc_detector *, int o, int s, int *dims, double *units,
                                  double *scale);
ORC_EXPORT int orc_filter(const orc_detector *, int idx, float *taps, int max_taps);
ORC_EXPORT int orc_num_candidates(const orc_detector *);
ORC_EXPORT void orc_candidates(const orc_detector *, orc_cand *out, int max);
ORC_EXPORT int orc_num_keypoints(const orc_detector *);
ORC_EXPORT void orc_keypoints(const orc_detector *, orc_key *out, int max);
ORC_EXPORT void orc_sort_by_strength(orc_detector *, int limit);
ORC_EXPORT int orc_describe(const orc_detector *, const orc_key *keys, int num, float *desc);
ORC_EXPORT int orc_blur(const float *src, float *dst, int nx, int ny, int nz, double ux, double uy,
                        double uz, const float *taps, int width);
ORC_EXPORT void orc_mesh(float *v /* [20][3][3] */, int32_t *idx /* [20][3] */);
ORC_EXPORT void orc_eigen3(const double *A9, double *Q9, double *L3);

#ifdef __cplusplus
}
#endif

#endif
