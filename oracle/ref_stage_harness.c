/* ref_stage_harness.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Stage-level access to the *unmodified* reference implementation. This file
 * does not contain reference code: it textually #includes the reference's
 * sift.c from where it lies under /root/reference at compile time (see
 * oracle/Makefile, target _ref/libsift3d_refstage.so) so that the reference's
 * `static` stages (set_im_SIFT3D, build_gpyr, build_dog, detect_extrema,
 * assign_orientations; sift.c:629-1167) and its private structs
 * (imtypes_private.h:73-223) become reachable, and exports small accessors
 * the Python tests use to dump pyramids, candidates, keypoints and
 * descriptors. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load the resulting library.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <imutil.h>
#include "sift.c" /* resolved via -I/root/reference/sift3d */

#define HX __attribute__((visibility("default")))

/* The fork has no public setter for voxel units (only read_nii sets them,
 * nifti.c:89-91); the tests need anisotropic volumes. */
HX void refstage_image_set_units(sift3d_image *im, double ux, double uy, double uz) {
    im->ux = ux;
    im->uy = uy;
    im->uz = uz;
}

/* Candidate record handed to Python (packed, 40 bytes). */
typedef struct {
    int32_t o, s;
    double x, y, z, sd;
    float strength;
    int32_t pad;
} refstage_cand;

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}

/* sift3d_detect_keypoints (sift.c:1217-1249) replayed stage by stage so the
 * candidate list can be captured between detect_extrema and
 * assign_orientations. times[5] = set_im, gpyr, dog, extrema, orient (s). */
HX int refstage_detect(sift3d_detector *det, const sift3d_image *im,
                       sift3d_keypoint_store *kp, refstage_cand *cand,
                       int max_cand, int *num_cand, double *times) {
    double t0, t1;
    int i, n;

    if (im->nc != 1)
        return SIFT3D_FAILURE;
    t0 = now_s();
    if (set_im_SIFT3D(det, im))
        return SIFT3D_FAILURE;
    t1 = now_s(); if (times) times[0] = t1 - t0; t0 = t1;
    if (build_gpyr(det))
        return SIFT3D_FAILURE;
    t1 = now_s(); if (times) times[1] = t1 - t0; t0 = t1;
    if (build_dog(det))
        return SIFT3D_FAILURE;
    t1 = now_s(); if (times) times[2] = t1 - t0; t0 = t1;
    if (detect_extrema(det, kp))
        return SIFT3D_FAILURE;
    t1 = now_s(); if (times) times[3] = t1 - t0; t0 = t1;

    n = (int) kp->slab.num;
    if (num_cand)
        *num_cand = n;
    if (cand != NULL) {
        for (i = 0; i < n && i < max_cand; i++) {
            const sift3d_keypoint *k = kp->buf + i;
            cand[i].o = k->o;
            cand[i].s = k->s;
            cand[i].x = k->xd;
            cand[i].y = k->yd;
            cand[i].z = k->zd;
            cand[i].sd = k->sd;
            cand[i].strength = k->strength;
            cand[i].pad = 0;
        }
    }

    if (assign_orientations(det, kp))
        return SIFT3D_FAILURE;
    t1 = now_s(); if (times) times[4] = t1 - t0;
    return SIFT3D_SUCCESS;
}

HX int refstage_num_octaves(const sift3d_detector *det) {
    return det->gpyr.num_octaves;
}

HX int refstage_num_levels(const sift3d_detector *det, int dog) {
    return dog ? det->dog.num_levels : det->gpyr.num_levels;
}

/* Pointer to a pyramid level's voxels plus its geometry. dog != 0 selects the
 * DoG pyramid. Returns NULL if (o, s) is out of range. */
HX const float *refstage_level(const sift3d_detector *det, int dog, int o, int s,
                               int *dims, double *units, double *scale) {
    const sift3d_pyramid *pyr = dog ? &det->dog : &det->gpyr;
    const sift3d_image *lv;

    if (pyr->levels == NULL || o < pyr->first_octave ||
        o > SIFT3D_PYR_LAST_OCTAVE(pyr) || s < pyr->first_level ||
        s > SIFT3D_PYR_LAST_LEVEL(pyr))
        return NULL;
    lv = SIFT3D_PYR_IM_GET(pyr, o, s);
    if (dims) { dims[0] = lv->nx; dims[1] = lv->ny; dims[2] = lv->nz; }
    if (units) { units[0] = lv->ux; units[1] = lv->uy; units[2] = lv->uz; }
    if (scale) *scale = lv->s;
    return lv->data;
}

/* Gaussian filter taps: idx -1 = first_gauss, 0..num_filters-1 = octave
 * filters (imtypes_private.h:98-103). Returns the width. */
HX int refstage_filter(const sift3d_detector *det, int idx, float *taps, int max_taps,
                       double *sigma) {
    const sift3d_gauss_filter *g;
    int i;

    if (idx < -1 || idx >= det->gss.num_filters)
        return -1;
    g = idx < 0 ? &det->gss.first_gauss : det->gss.gauss_octave + idx;
    if (sigma) *sigma = g->sigma;
    for (i = 0; i < g->f.width && i < max_taps; i++)
        taps[i] = g->f.kernel[i];
    return g->f.width;
}

/* The icosahedron table as the reference builds it (sift.c:148-259). */
HX void refstage_mesh(const sift3d_detector *det, float *v /* [20][3][3] */,
                      int32_t *idx /* [20][3] */) {
    int i, j;
    for (i = 0; i < ICOS_NFACES; i++) {
        for (j = 0; j < 3; j++) {
            v[(i * 3 + j) * 3 + 0] = det->mesh.tri[i].v[j].x;
            v[(i * 3 + j) * 3 + 1] = det->mesh.tri[i].v[j].y;
            v[(i * 3 + j) * 3 + 2] = det->mesh.tri[i].v[j].z;
            idx[i * 3 + j] = det->mesh.tri[i].idx[j];
        }
    }
}

/* Keypoint record handed to Python (packed, 88 bytes). */
typedef struct {
    int32_t o, s;
    double x, y, z, sd;
    float strength;
    float R[9];
    int32_t pad;
} refstage_key;

HX int refstage_num_keypoints(const sift3d_keypoint_store *kp) {
    return (int) kp->slab.num;
}

HX void refstage_keypoints(const sift3d_keypoint_store *kp, refstage_key *out, int max) {
    int i, j;
    for (i = 0; i < (int) kp->slab.num && i < max; i++) {
        const sift3d_keypoint *k = kp->buf + i;
        out[i].o = k->o;
        out[i].s = k->s;
        out[i].x = k->xd;
        out[i].y = k->yd;
        out[i].z = k->zd;
        out[i].sd = k->sd;
        out[i].strength = k->strength;
        for (j = 0; j < 9; j++)
            out[i].R[j] = k->r_data[j];
        out[i].pad = 0;
    }
}

/* Overwrite a keypoint store with caller-provided keypoints (used to run the
 * reference's descriptor stage on an arbitrary keypoint list). */
HX int refstage_set_keypoints(sift3d_keypoint_store *kp, const refstage_key *in, int num) {
    int i, j;
    if (resize_Keypoint_store(kp, num))
        return SIFT3D_FAILURE;
    for (i = 0; i < num; i++) {
        sift3d_keypoint *k = kp->buf + i;
        if (init_Keypoint(k))
            return SIFT3D_FAILURE;
        k->o = in[i].o;
        k->s = in[i].s;
        k->xd = in[i].x;
        k->yd = in[i].y;
        k->zd = in[i].z;
        k->sd = in[i].sd;
        k->strength = in[i].strength;
        for (j = 0; j < 9; j++)
            k->r_data[j] = in[i].R[j];
    }
    return SIFT3D_SUCCESS;
}

/* One separable blur through the reference's apply_Sep_FIR_filter
 * (imutil.c:1127-1206) with caller-supplied taps, dims and voxel units;
 * unit = 1.0 as build_gpyr passes it (sift.c:675). */
HX int refstage_blur(const float *src, float *dst, int nx, int ny, int nz,
                     double ux, double uy, double uz, const float *taps, int width) {
    sift3d_image a, b;
    sift3d_sep_fir_filter f;
    int ret;

    if (init_im_with_dims(&a, nx, ny, nz, 1))
        return SIFT3D_FAILURE;
    init_im(&b);
    a.ux = ux; a.uy = uy; a.uz = uz;
    memcpy(a.data, src, (size_t) nx * ny * nz * sizeof(float));
    f.kernel = (float *) taps;
    f.dim = 3;
    f.width = width;
    f.symmetric = 1;
    ret = apply_Sep_FIR_filter(&a, &b, &f, 1.0);
    if (!ret)
        memcpy(dst, b.data, (size_t) nx * ny * nz * sizeof(float));
    im_free(&a);
    im_free(&b);
    return ret;
}

/* The reference's eigen-decomposition (imutil.c:984-1067, LAPACK dsyevd) of a
 * symmetric 3x3: L ascending, Q columns = eigenvectors. */
HX int refstage_eigen3(const double *A9, double *Q9, double *L3) {
    sift3d_mat_rm A, Q, L;
    int ret, i;

    if (init_Mat_rm(&A, 3, 3, SIFT3D_DOUBLE, SIFT3D_TRUE) ||
        init_Mat_rm(&Q, 0, 0, SIFT3D_DOUBLE, SIFT3D_TRUE) ||
        init_Mat_rm(&L, 0, 0, SIFT3D_DOUBLE, SIFT3D_TRUE))
        return SIFT3D_FAILURE;
    for (i = 0; i < 9; i++)
        A.u.data_double[i] = A9[i];
    ret = eigen_Mat_rm(&A, &Q, &L);
    if (!ret) {
        for (i = 0; i < 9; i++) Q9[i] = Q.u.data_double[i];
        for (i = 0; i < 3; i++) L3[i] = L.u.data_double[i];
    }
    cleanup_Mat_rm(&A);
    cleanup_Mat_rm(&Q);
    cleanup_Mat_rm(&L);
    return ret;
}

HX size_t refstage_sizeof(int what) {
    switch (what) {
    case 0: return sizeof(sift3d_keypoint);
    case 1: return sizeof(sift3d_descriptor);
    case 2: return sizeof(sift3d_image);
    case 3: return sizeof(sift3d_detector);
    default: return 0;
    }
}
