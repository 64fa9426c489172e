/* sift3d_oracle.c -- CPU restatement ("port") of the reference's detect + describe
 * path, written from the specification in SURVEY.md Appendix A and the cited
 * reference lines. TEST INFRASTRUCTURE ONLY: it may be loaded by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product,
 * and it is never the thing measured as the product.
 *
 * Parity status: PINNED against the reference itself -- tests/test_oracle.py checks
 * this restatement against (a) oracle/_ref (the unmodified reference compiled from
 * /root/reference by oracle/Makefile) whenever that library is present and (b) the
 * golden vectors under tests/golden/ that tests/golden/make_golden.py generated
 * from oracle/_ref. The reference itself ships no tests or fixtures.
 * The only arithmetic not taken from the reference is the symmetric 3x3
 * eigen-decomposition: the reference calls LAPACK dsyevd_ (imutil.c:1027-1045, not
 * vendored, no pinned version); a cyclic Jacobi iteration in double stands in here.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (no FMA contraction, like the
 * reference's x86-64 build), see oracle/Makefile. Citations below are to
 * /root/reference/sift3d/.
 */
#include "sift3d_oracle.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MAX_TAPS 513

typedef struct {
    float *data;
    int nx, ny, nz;
    double ux, uy, uz, s;
} level_t;

typedef struct {
    float k[MAX_TAPS];
    int width;
} filt_t;

struct orc_detector {
    double peak_thresh, corner_thresh, sigma_n, sigma0;
    int num_kp_levels;
    /* pyramid */
    int num_octaves, num_levels; /* Gaussian levels per octave, s = -1 .. num_levels-2 */
    int pnx, pny, pnz;           /* dims the pyramid was sized for */
    double punits[3];            /* units captured at that time (sift.c:651-656) */
    level_t *levels;
    filt_t first, *octave;       /* octave[i]: level i-1 -> i */
    /* results */
    orc_cand *cand;
    int ncand;
    orc_key *keys;
    int nkeys;
};

static level_t *LV(const orc_detector *d, int o, int s) {
    return d->levels + (size_t) o * d->num_levels + (s + 1);
}

/* ---- Gaussian filters: init_Gauss_filter imutil.c:1267-1319,
 *      init_Gauss_incremental_filter imutil.c:1322-1343 ---- */
static int gauss_taps(double sigma, filt_t *f) {
    int hw = sigma > 0 ? (int) ceil(sigma * 3.0) : 1;
    float acc = 0;
    int i;
    if (hw < 1) hw = 1;
    f->width = 2 * hw + 1;
    if (f->width > MAX_TAPS) return -1;
    for (i = 0; i < f->width; i++) {
        double x = (double) i - hw;
        x /= sigma + DBL_EPSILON;
        f->k[i] = (float) exp(-0.5 * x * x);
        acc += f->k[i];
    }
    for (i = 0; i < f->width; i++) f->k[i] /= acc;
    return 0;
}

static int gauss_incremental(double s_cur, double s_next, filt_t *f) {
    if (s_cur > s_next) return -1;
    return gauss_taps(sqrt(s_next * s_next - s_cur * s_cur), f);
}

/* ---- one line of convolve_sep_gen (imutil.c:742-861): interior [uhw, n-2-uhw] with the
 *      sampling coordinate carried across taps, mirrored boundaries elsewhere ---- */
static void conv_line(const float *src, float *dst, long stride, int n, const float *k, int hw,
                      float uf) {
    const int uhw = (int) ceilf(hw * uf);          /* imutil.c:756-757 */
    const int start = uhw, end = n - 2 - uhw;      /* imutil.c:758-763 */
    const int dim_end = n - 1;                     /* imutil.c:753 */
    const float conv_eps = 0.1f;                   /* imutil.c:752 */
    int p, d;
    for (p = 0; p < n; p++) {
        float acc = 0.0f; /* im_zero, imutil.c:777 */
        if (p >= start && p <= end) {
            float c = (float) p;
            for (d = -hw; d <= hw; d++) {
                const float step = d * uf;
                int lo;
                float fr;
                c -= step;
                lo = (int) c;
                fr = c - (float) lo;
                acc += k[d + hw] * ((1.0f - fr) * src[(long) lo * stride] +
                                    fr * src[(long) (lo + 1) * stride]);
                c += step;
            }
        } else {
            for (d = -hw; d <= hw; d++) {
                const float step = d * uf;
                float c = (float) p;
                float fr;
                int lo;
                c -= step;
                if ((int) c < 0) {
                    c = -c;
                } else if ((int) c >= dim_end) {
                    c = 2.0f * dim_end - c - conv_eps;
                }
                lo = (int) c;
                fr = c - (float) lo;
                /* lines shorter than the kernel make the reference read out of bounds
                 * here (single mirror); keep the access legal instead */
                if (lo > n - 2) lo = n - 2;
                if (lo < 0) lo = 0;
                acc += k[d + hw] * ((1.0f - fr) * src[(long) lo * stride] +
                                    fr * src[(long) (lo + 1) * stride]);
            }
        }
        dst[(long) p * stride] = acc;
    }
}

/* apply_Sep_FIR_filter (imutil.c:1127-1206) with unit = 1.0 (sift.c:675): x, y, z passes,
 * each rounded to fp32. The reference permutes the volume so that the filtered axis is x;
 * walking the axis with a stride visits the same samples in the same order. */
static int blur3(const float *src, float *dst, int nx, int ny, int nz, const double u[3],
                 const filt_t *f) {
    const size_t n = (size_t) nx * ny * nz;
    const int hw = f->width / 2;
    const float ufx = (float) (1.0 / u[0]), ufy = (float) (1.0 / u[1]), ufz = (float) (1.0 / u[2]);
    float *t1 = (float *) malloc(n * sizeof(float));
    float *t2 = (float *) malloc(n * sizeof(float));
    long i;
    if (!t1 || !t2) { free(t1); free(t2); return -1; }
#pragma omp parallel for schedule(static)
    for (i = 0; i < (long) ny * nz; i++)
        conv_line(src + (size_t) i * nx, t1 + (size_t) i * nx, 1, nx, f->k, hw, ufx);
#pragma omp parallel for schedule(static)
    for (i = 0; i < (long) nx * nz; i++) {
        const size_t base = (size_t) (i / nx) * nx * ny + (size_t) (i % nx);
        conv_line(t1 + base, t2 + base, nx, ny, f->k, hw, ufy);
    }
#pragma omp parallel for schedule(static)
    for (i = 0; i < (long) nx * ny; i++)
        conv_line(t2 + i, dst + i, (long) nx * ny, nz, f->k, hw, ufz);
    free(t1);
    free(t2);
    return 0;
}

int orc_blur(const float *src, float *dst, int nx, int ny, int nz, double ux, double uy,
             double uz, const float *taps, int width) {
    filt_t f;
    const double u[3] = {ux, uy, uz};
    if (width > MAX_TAPS || !(width & 1)) return -1;
    memcpy(f.k, taps, sizeof(float) * (size_t) width);
    f.width = width;
    return blur3(src, dst, nx, ny, nz, u, &f);
}

/* ---- detector lifecycle ---- */
orc_detector *orc_make(void) {
    orc_detector *d = (orc_detector *) calloc(1, sizeof(*d));
    if (!d) return NULL;
    d->peak_thresh = 0.1;   /* sift.c:31-35 */
    d->num_kp_levels = 3;
    d->corner_thresh = 0.4;
    d->sigma_n = 1.15;
    d->sigma0 = 1.6;
    return d;
}

static void free_pyramid(orc_detector *d) {
    int i;
    if (d->levels)
        for (i = 0; i < d->num_octaves * d->num_levels; i++) free(d->levels[i].data);
    free(d->levels);
    free(d->octave);
    d->levels = NULL;
    d->octave = NULL;
    d->num_octaves = 0;
}

void orc_free(orc_detector *d) {
    if (!d) return;
    free_pyramid(d);
    free(d->cand);
    free(d->keys);
    free(d);
}

int orc_set_params(orc_detector *d, double peak, double corner, int kp, double sigma_n,
                   double sigma0) {
    /* ranges: sift.c:501, 515, 542, 558 */
    if (peak <= 0.0 || peak > 1 || corner < 0.0 || corner > 1.0 || sigma_n < 0.0 || sigma0 < 0.0 ||
        kp < 1)
        return -1;
    d->peak_thresh = peak;
    d->corner_thresh = corner;
    if (kp != d->num_kp_levels || sigma_n != d->sigma_n || sigma0 != d->sigma0) free_pyramid(d);
    d->num_kp_levels = kp;
    d->sigma_n = sigma_n;
    d->sigma0 = sigma0;
    return 0;
}

/* resize_SIFT3D sift.c:427-475, resize_Pyramid imutil.c:1464-1554,
 * set_scales_Pyramid imutil.c:1564-1599, make_gss imutil.c:1360-1409 */
static int plan(orc_detector *d, int nx, int ny, int nz, const double u[3]) {
    int mind = nx < ny ? nx : ny, last_octave, o, s, dx = nx, dy = ny, dz = nz;
    double f = 1.0;
    mind = mind < nz ? mind : nz;
    last_octave = (int) log2((double) mind) - 3;
    if (last_octave < 0) {
        fprintf(stderr, "oracle: input image is too small: must have at least 8 voxels in each "
                        "dimension\n");
        return -1;
    }
    free_pyramid(d);
    d->num_octaves = last_octave + 1;
    d->num_levels = d->num_kp_levels + 3;
    d->levels = (level_t *) calloc((size_t) d->num_octaves * d->num_levels, sizeof(level_t));
    d->octave = (filt_t *) calloc((size_t) d->num_levels - 1, sizeof(filt_t));
    if (!d->levels || !d->octave) return -1;
    for (o = 0; o < d->num_octaves; o++) {
        for (s = -1; s < d->num_levels - 1; s++) {
            level_t *L = LV(d, o, s);
            L->nx = dx; L->ny = dy; L->nz = dz;
            L->ux = u[0] * f; L->uy = u[1] * f; L->uz = u[2] * f;
            L->s = d->sigma0 * pow(2.0, o + (double) s / d->num_kp_levels);
            L->data = (float *) malloc(sizeof(float) * (size_t) dx * dy * dz);
            if (!L->data) return -1;
        }
        dx /= 2; dy /= 2; dz /= 2;   /* imutil.c:1545-1548 */
        f *= 2;
    }
    if (LV(d, 0, -1)->s < d->sigma_n) return -1;   /* imutil.c:1582-1588 */
    if (gauss_incremental(d->sigma_n, LV(d, 0, -1)->s, &d->first)) return -1;
    for (s = -1; s < d->num_levels - 2; s++)
        if (gauss_incremental(LV(d, 0, s)->s, LV(d, 0, s + 1)->s, &d->octave[s + 1])) return -1;
    d->pnx = nx; d->pny = ny; d->pnz = nz;
    d->punits[0] = u[0]; d->punits[1] = u[1]; d->punits[2] = u[2];
    return 0;
}

/* ---- symmetric 3x3 eigen-decomposition (stands in for eigen_Mat_rm, imutil.c:984-1067):
 *      eigenvalues ascending, eigenvectors in the columns of V ---- */
static void eig3(const double Ain[3][3], double V[3][3], double w[3]) {
    double A[3][3];
    int i, j, sweep, pq, k;
    for (i = 0; i < 3; i++)
        for (j = 0; j < 3; j++) { A[i][j] = Ain[i][j]; V[i][j] = i == j; }
    for (sweep = 0; sweep < 64; sweep++) {
        if (fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]) == 0.0) break;
        for (pq = 0; pq < 3; pq++) {
            const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2, r = 3 - p - q;
            const double apq = A[p][q], app = A[p][p], aqq = A[q][q];
            double theta, t, c, s, tau, arp, arq;
            if (apq == 0.0) continue;
            if (fabs(apq) < DBL_MIN ||
                (fabs(app) + fabs(apq) == fabs(app) && fabs(aqq) + fabs(apq) == fabs(aqq))) {
                A[p][q] = A[q][p] = 0.0;
                continue;
            }
            theta = (aqq - app) / (2.0 * apq);
            t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            c = 1.0 / sqrt(t * t + 1.0);
            s = t * c;
            tau = s / (1.0 + c);
            A[p][p] = app - t * apq;
            A[q][q] = aqq + t * apq;
            A[p][q] = A[q][p] = 0.0;
            arp = A[r][p]; arq = A[r][q];
            A[r][p] = A[p][r] = arp - s * (arq + tau * arp);
            A[r][q] = A[q][r] = arq + s * (arp - tau * arq);
            for (k = 0; k < 3; k++) {
                const double vkp = V[k][p], vkq = V[k][q];
                V[k][p] = vkp - s * (vkq + tau * vkp);
                V[k][q] = vkq + s * (vkp - tau * vkq);
            }
        }
    }
    for (i = 0; i < 3; i++) w[i] = A[i][i];
    for (i = 0; i < 3; i++) { /* ascending order, columns follow */
        const int a = i == 1 ? 1 : 0, b = i == 1 ? 2 : 1;
        if (w[a] > w[b]) {
            double tw = w[a]; w[a] = w[b]; w[b] = tw;
            for (k = 0; k < 3; k++) { double tv = V[k][a]; V[k][a] = V[k][b]; V[k][b] = tv; }
        }
    }
}

void orc_eigen3(const double *A9, double *Q9, double *L3) {
    double A[3][3], V[3][3];
    int i, j;
    for (i = 0; i < 3; i++) for (j = 0; j < 3; j++) A[i][j] = A9[i * 3 + j];
    eig3(A, V, L3);
    for (i = 0; i < 3; i++) for (j = 0; j < 3; j++) Q9[i * 3 + j] = V[i][j];
}

/* ---- window bounds of IM_LOOP_SPHERE_START (sift.c:86-99) ---- */
static void window(float c, double rad_over_u_is_double, double rad, float uf, int n, int *lo,
                   int *hi) {
    float a, b;
    if (rad_over_u_is_double) { /* rad is a double (orientation) */
        a = floorf((float) ((double) c - rad / (double) uf));
        b = ceilf((float) ((double) c + rad / (double) uf));
    } else {                    /* rad is a float (descriptor) */
        const float r = (float) rad;
        a = floorf(c - r / uf);
        b = ceilf(c + r / uf);
    }
    *lo = (int) (a > 1 ? a : 1);
    *hi = (int) (b < (float) (n - 2) ? b : (float) (n - 2));
}

#define VOX(L, x, y, z) ((L)->data[(size_t) (x) + (size_t) (y) * (L)->nx + (size_t) (z) * (L)->nx * (L)->ny])

/* IM_GET_GRAD_ISO (sift.c:140-145, immacros.h:105-111) */
static void grad_iso(const level_t *L, int x, int y, int z, float g[3]) {
    g[0] = 0.5f * (VOX(L, x + 1, y, z) - VOX(L, x - 1, y, z));
    g[1] = 0.5f * (VOX(L, x, y + 1, z) - VOX(L, x, y - 1, z));
    g[2] = 0.5f * (VOX(L, x, y, z + 1) - VOX(L, x, y, z - 1));
    g[0] *= 1.0f / (float) L->ux;
    g[1] *= 1.0f / (float) L->uy;
    g[2] *= 1.0f / (float) L->uz;
}

/* assign_eig_ori + assign_orientation_thresh (sift.c:926-1102). Returns 1 if kept. */
static int orient(const level_t *L, float cx, float cy, float cz, double sigma,
                  double corner_thresh, float R[9]) {
    const double rad = sigma * 3.0;
    const float uxf = (float) L->ux, uyf = (float) L->uy, uzf = (float) L->uz;
    double A[3][3] = {{0}}, V[3][3], w[3], corner = DBL_MAX;
    float gw[3] = {0, 0, 0}, v[2][3];
    int x0, x1, y0, y1, z0, z1, x, y, z, i;
    window(cx, 1, rad, uxf, L->nx, &x0, &x1);
    window(cy, 1, rad, uyf, L->ny, &y0, &y1);
    window(cz, 1, rad, uzf, L->nz, &z0, &z1);
    for (z = z0; z <= z1; z++)
        for (y = y0; y <= y1; y++)
            for (x = x0; x <= x1; x++) {
                const float dx = ((float) x - cx) * uxf, dy = ((float) y - cy) * uyf,
                            dz = ((float) z - cz) * uzf;
                const float sq = dx * dx + dy * dy + dz * dz;
                float g[3], wt;
                if (sq > rad * rad) continue;
                wt = expf(-0.5 * sq / (sigma * sigma));   /* sift.c:972 */
                grad_iso(L, x, y, z, g);
                A[0][0] += (double) g[0] * g[0] * wt;
                A[0][1] += (double) g[0] * g[1] * wt;
                A[0][2] += (double) g[0] * g[2] * wt;
                A[1][1] += (double) g[1] * g[1] * wt;
                A[1][2] += (double) g[1] * g[2] * wt;
                A[2][2] += (double) g[2] * g[2] * wt;
                gw[0] = gw[0] + g[0] * wt;
                gw[1] = gw[1] + g[1] * wt;
                gw[2] = gw[2] + g[2] * wt;
            }
    A[1][0] = A[0][1]; A[2][0] = A[0][2]; A[2][1] = A[1][2];
    if (gw[0] * gw[0] + gw[1] * gw[1] + gw[2] * gw[2] < (float) 1E-10) return 0;  /* sift.c:997 */
    eig3(A, V, w);
    for (i = 0; i < 2; i++)
        if (fabs(w[i] / w[i + 1]) > 0.90) return 0;                                /* sift.c:1011-1015 */
    for (i = 0; i < 2; i++) {
        const int e = 2 - i;
        float vr[3] = {(float) V[0][e], (float) V[1][e], (float) V[2][e]};
        const double d = gw[0] * vr[0] + gw[1] * vr[1] + gw[2] * vr[2];
        const double cosang = d / (sqrtf(vr[0] * vr[0] + vr[1] * vr[1] + vr[2] * vr[2]) *
                                   sqrtf(gw[0] * gw[0] + gw[1] * gw[1] + gw[2] * gw[2]));
        const float sgn = d > 0.0 ? 1.0f : -1.0f;
        if (fabs(cosang) < corner) corner = fabs(cosang);
        vr[0] *= sgn; vr[1] *= sgn; vr[2] *= sgn;
        R[0 + i] = vr[0]; R[3 + i] = vr[1]; R[6 + i] = vr[2];
        v[i][0] = vr[0]; v[i][1] = vr[1]; v[i][2] = vr[2];
    }
    R[2] = v[0][1] * v[1][2] - v[0][2] * v[1][1];
    R[5] = v[0][2] * v[1][0] - v[0][0] * v[1][2];
    R[8] = v[0][0] * v[1][1] - v[0][1] * v[1][0];
    return corner < corner_thresh ? 0 : 1;
}

int orc_detect(orc_detector *d, const float *vol, int nx, int ny, int nz, double ux, double uy,
               double uz) {
    const double u0[3] = {ux, uy, uz};
    const size_t n0 = (size_t) nx * ny * nz;
    const int kp = d->num_kp_levels;
    float *im, mx = 0.0f;
    size_t i;
    int o, s, cap = 0, j;

    /* set_im_SIFT3D (sift.c:629-659): re-plan only when the dims change */
    if (!d->levels || d->pnx != nx || d->pny != ny || d->pnz != nz)
        if (plan(d, nx, ny, nz, u0)) return -1;
    /* im_scale (imutil.c:681-713) */
    if (!(im = (float *) malloc(n0 * sizeof(float)))) return -1;
    for (i = 0; i < n0; i++) { const float a = fabsf(vol[i]); mx = mx > a ? mx : a; }
    if (mx != 0.0f) for (i = 0; i < n0; i++) im[i] = vol[i] / mx;
    else memcpy(im, vol, n0 * sizeof(float));

    /* units: octave 0 follows the image, octaves >= 1 keep the planned units
     * (imutil.c:1145, 596-603; sift.c:651-656) */
    for (s = -1; s < d->num_levels - 1; s++) { LV(d, 0, s)->ux = ux; LV(d, 0, s)->uy = uy; LV(d, 0, s)->uz = uz; }

    /* build_gpyr (sift.c:662-711) */
    {
        level_t *L = LV(d, 0, -1);
        if (blur3(im, L->data, L->nx, L->ny, L->nz, u0, &d->first)) { free(im); return -1; }
    }
    free(im);
    for (o = 0; o < d->num_octaves; o++) {
        for (s = 0; s < d->num_levels - 1; s++) {
            level_t *P = LV(d, o, s - 1), *L = LV(d, o, s);
            const double u[3] = {P->ux, P->uy, P->uz};
            if (blur3(P->data, L->data, L->nx, L->ny, L->nz, u, &d->octave[s])) return -1;
        }
        if (o != d->num_octaves - 1) {  /* im_downsample_2x, imutil.c:591-617 */
            const int ds = d->num_levels - 2 - 2 > -1 ? d->num_levels - 2 - 2 : -1;
            const level_t *S = LV(d, o, ds);
            level_t *T = LV(d, o + 1, -1);
            int x, y, z;
            for (z = 0; z < T->nz; z++)
                for (y = 0; y < T->ny; y++)
                    for (x = 0; x < T->nx; x++) VOX(T, x, y, z) = VOX(S, 2 * x, 2 * y, 2 * z);
        }
    }

    /* build_dog + detect_extrema (sift.c:713-871); D_s = G_s - G_{s+1} computed on the fly */
    d->ncand = 0;
    for (o = 0; o < d->num_octaves; o++)
        for (s = 0; s < kp; s++) {
            const level_t *G0 = LV(d, o, s - 1), *G1 = LV(d, o, s), *G2 = LV(d, o, s + 1),
                          *G3 = LV(d, o, s + 2);
            const int lx = G1->nx, ly = G1->ny, lz = G1->nz;
            const size_t nv = (size_t) lx * ly * lz;
            float dogmax = 0.0f, thr;
            int x, y, z;
            for (i = 0; i < nv; i++) { const float a = fabsf(G1->data[i] - G2->data[i]); dogmax = dogmax > a ? dogmax : a; }
            thr = (float) (d->peak_thresh * dogmax);
#define DOG(x, y, z) (VOX(G1, x, y, z) - VOX(G2, x, y, z))
            for (z = 1; z <= lz - 2; z++)
                for (y = 1; y <= ly - 2; y++)
                    for (x = 1; x <= lx - 2; x++) {
                        const float v = DOG(x, y, z);
                        float nb[8];
                        int all_gt = 1, all_lt = 1;
                        if (!(v > thr || v < -thr)) continue;
                        nb[0] = VOX(G0, x, y, z) - VOX(G1, x, y, z);
                        nb[1] = VOX(G2, x, y, z) - VOX(G3, x, y, z);
                        nb[2] = DOG(x + 1, y, z); nb[3] = DOG(x - 1, y, z);
                        nb[4] = DOG(x, y + 1, z); nb[5] = DOG(x, y - 1, z);
                        nb[6] = DOG(x, y, z - 1); nb[7] = DOG(x, y, z + 1);
                        for (j = 0; j < 8; j++) { all_gt &= v > nb[j]; all_lt &= v < nb[j]; }
                        if (!all_gt && !all_lt) continue;
                        if (d->ncand == cap) {
                            cap = cap ? cap * 2 : 1024;
                            d->cand = (orc_cand *) realloc(d->cand, sizeof(orc_cand) * (size_t) cap);
                            if (!d->cand) return -1;
                        }
                        d->cand[d->ncand].o = o; d->cand[d->ncand].s = s;
                        d->cand[d->ncand].x = x; d->cand[d->ncand].y = y; d->cand[d->ncand].z = z;
                        d->cand[d->ncand].sd = G1->s;
                        d->cand[d->ncand].strength = fabsf(v);
                        d->cand[d->ncand].pad = 0;
                        d->ncand++;
                    }
#undef DOG
        }

    /* assign_orientations (sift.c:1109-1167) */
    free(d->keys);
    d->keys = (orc_key *) calloc((size_t) (d->ncand > 0 ? d->ncand : 1), sizeof(orc_key));
    if (!d->keys) return -1;
    {
        char *keep = (char *) calloc((size_t) d->ncand + 1, 1);
        orc_key *tmp = (orc_key *) calloc((size_t) d->ncand + 1, sizeof(orc_key));
        if (!keep || !tmp) { free(keep); free(tmp); return -1; }
#pragma omp parallel for schedule(dynamic, 8)
        for (j = 0; j < d->ncand; j++) {
            const orc_cand *c = d->cand + j;
            tmp[j].o = c->o; tmp[j].s = c->s; tmp[j].x = c->x; tmp[j].y = c->y; tmp[j].z = c->z;
            tmp[j].sd = c->sd; tmp[j].strength = c->strength;
            keep[j] = (char) orient(LV(d, c->o, c->s), (float) c->x, (float) c->y, (float) c->z,
                                    1.5 * c->sd, d->corner_thresh, tmp[j].R);
        }
        /* in-order compaction; copy_Keypoint leaves `strength` behind (sift.c:372-384), so
         * survivor k carries the strength of candidate k */
        d->nkeys = 0;
        for (j = 0; j < d->ncand; j++) {
            if (!keep[j]) continue;
            d->keys[d->nkeys] = tmp[j];
            d->keys[d->nkeys].strength = d->cand[d->nkeys].strength;
            d->nkeys++;
        }
        free(keep);
        free(tmp);
    }
    return 0;
}

int orc_num_octaves(const orc_detector *d) { return d->num_octaves; }
int orc_num_levels(const orc_detector *d) { return d->num_levels; }

const float *orc_level(const orc_detector *d, int o, int s, int *dims, double *units, double *scale) {
    const level_t *L;
    if (!d->levels || o < 0 || o >= d->num_octaves || s < -1 || s > d->num_levels - 2) return NULL;
    L = LV(d, o, s);
    if (dims) { dims[0] = L->nx; dims[1] = L->ny; dims[2] = L->nz; }
    if (units) { units[0] = L->ux; units[1] = L->uy; units[2] = L->uz; }
    if (scale) *scale = L->s;
    return L->data;
}

int orc_filter(const orc_detector *d, int idx, float *taps, int max_taps) {
    const filt_t *f;
    int i;
    if (!d->octave || idx < -1 || idx >= d->num_levels - 1) return -1;
    f = idx < 0 ? &d->first : &d->octave[idx];
    for (i = 0; i < f->width && i < max_taps; i++) taps[i] = f->k[i];
    return f->width;
}

int orc_num_candidates(const orc_detector *d) { return d->ncand; }
void orc_candidates(const orc_detector *d, orc_cand *out, int max) {
    memcpy(out, d->cand, sizeof(orc_cand) * (size_t) (d->ncand < max ? d->ncand : max));
}
int orc_num_keypoints(const orc_detector *d) { return d->nkeys; }
void orc_keypoints(const orc_detector *d, orc_key *out, int max) {
    memcpy(out, d->keys, sizeof(orc_key) * (size_t) (d->nkeys < max ? d->nkeys : max));
}

/* keypoint_strength_cmp + sort_by_strength (sift.c:1832-1837, 1885-1900) */
static int strength_cmp(const void *a, const void *b) {
    return (((const orc_key *) a)->strength < ((const orc_key *) b)->strength) ? 1 : -1;
}
void orc_sort_by_strength(orc_detector *d, int limit) {
    qsort(d->keys, (size_t) d->nkeys, sizeof(orc_key), strength_cmp);
    if (limit != 0 && d->nkeys > limit) d->nkeys = limit;
}

/* ---- icosahedron (init_geometry, sift.c:148-259) ---- */
typedef struct { float v[3][3]; int idx[3]; } tri_t;
static tri_t g_mesh[20];
static int g_mesh_ready = 0;

static void build_mesh(void) {
    const double gr = 1.6180339887;  /* sift.c:48 */
    const float vert[12][3] = {{0, 1, gr}, {0, -1, gr}, {0, 1, -gr}, {0, -1, -gr}, {1, gr, 0},
                               {-1, gr, 0}, {1, -gr, 0}, {-1, -gr, 0}, {gr, 0, 1}, {-gr, 0, 1},
                               {gr, 0, -1}, {-gr, 0, -1}};
    const int faces[20][3] = {{0, 1, 8}, {0, 8, 4}, {0, 4, 5}, {0, 5, 9}, {0, 9, 1}, {1, 6, 8},
                              {8, 6, 10}, {8, 10, 4}, {4, 10, 2}, {4, 2, 5}, {5, 2, 11},
                              {5, 11, 9}, {9, 11, 7}, {9, 7, 1}, {1, 7, 6}, {3, 6, 7}, {3, 7, 11},
                              {3, 11, 2}, {3, 2, 10}, {3, 10, 6}};
    int i, j, k;
    for (i = 0; i < 20; i++) {
        tri_t *t = &g_mesh[i];
        float a[3], b[3], n[3];
        for (j = 0; j < 3; j++) {
            const float *p = vert[faces[i][j]];
            const float mag = sqrtf(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
            t->idx[j] = faces[i][j];
            /* SIFT3D_CVEC_SCALE(v, 1.0f / mag) expands to v * 1.0f / mag (sift.c:228) */
            for (k = 0; k < 3; k++) t->v[j][k] = p[k] * 1.0f / mag;
        }
        for (k = 0; k < 3; k++) { a[k] = t->v[2][k] - t->v[1][k]; b[k] = t->v[1][k] - t->v[0][k]; }
        n[0] = a[1] * b[2] - a[2] * b[1];
        n[1] = a[2] * b[0] - a[0] * b[2];
        n[2] = a[0] * b[1] - a[1] * b[0];
        if (n[0] * t->v[0][0] + n[1] * t->v[0][1] + n[2] * t->v[0][2] < 0)  /* sift.c:237-242 */
            for (k = 0; k < 3; k++) { const float tmp = t->v[0][k]; t->v[0][k] = t->v[1][k]; t->v[1][k] = tmp; }
    }
    g_mesh_ready = 1;
}

void orc_mesh(float *v, int32_t *idx) {
    int i, j, k;
    if (!g_mesh_ready) build_mesh();
    for (i = 0; i < 20; i++)
        for (j = 0; j < 3; j++) {
            for (k = 0; k < 3; k++) v[(i * 3 + j) * 3 + k] = g_mesh[i].v[j][k];
            idx[i * 3 + j] = g_mesh[i].idx[j];
        }
}

static const double bary_eps = FLT_EPSILON * 1E1;  /* sift.c:40 */

/* cart2bary (sift.c:268-297) */
static int cart2bary(const float c[3], const tri_t *t, float bary[3], float *kout) {
    float e1[3], e2[3], tv[3], p[3], q[3], det, det_inv;
    int i;
    for (i = 0; i < 3; i++) { e1[i] = t->v[1][i] - t->v[0][i]; e2[i] = t->v[2][i] - t->v[0][i]; }
    p[0] = c[1] * e2[2] - c[2] * e2[1];
    p[1] = c[2] * e2[0] - c[0] * e2[2];
    p[2] = c[0] * e2[1] - c[1] * e2[0];
    det = e1[0] * p[0] + e1[1] * p[1] + e1[2] * p[2];
    if (fabsf(det) < bary_eps) return -1;
    det_inv = 1.0f / det;
    for (i = 0; i < 3; i++) tv[i] = t->v[0][i] * -1.0f;
    q[0] = tv[1] * e1[2] - tv[2] * e1[1];
    q[1] = tv[2] * e1[0] - tv[0] * e1[2];
    q[2] = tv[0] * e1[1] - tv[1] * e1[0];
    bary[1] = det_inv * (tv[0] * p[0] + tv[1] * p[1] + tv[2] * p[2]);
    bary[2] = det_inv * (c[0] * q[0] + c[1] * q[1] + c[2] * q[2]);
    bary[0] = 1.0f - bary[1] - bary[2];
    *kout = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * det_inv;
    return 0;
}

/* extract_descrip (sift.c:1442-1536) incl. SIFT3D_desc_acc_interp (sift.c:1295-1399),
 * icos_hist_bin (sift.c:1254-1291) and normalize_desc (sift.c:1402-1429) */
static void describe_one(const level_t *L, const orc_key *key, float *hist) {
    const float sigma = key->sd * 7.071067812;
    const float rad = 2.0 * sigma;
    const float half_w = rad / sqrt(2);
    const float desc_w = 2.0f * half_w;
    const float hist_w = desc_w / 4;
    const float bin_fctr = 1.0f / hist_w;
    const float uxf = (float) L->ux, uyf = (float) L->uy, uzf = (float) L->uz;
    const float cx = key->x, cy = key->y, cz = key->z;
    const float *R = key->R;
    const float trunc = 0.2f * 128.0f / ORC_DESC_NUMEL;  /* sift.c:45 */
    int x0, x1, y0, y1, z0, z1, x, y, z, i, pass;
    memset(hist, 0, sizeof(float) * ORC_DESC_NUMEL);
    window(cx, 0, rad, uxf, L->nx, &x0, &x1);
    window(cy, 0, rad, uyf, L->ny, &y0, &y1);
    window(cz, 0, rad, uzf, L->nz, &z0, &z1);
    for (z = z0; z <= z1; z++)
        for (y = y0; y <= y1; y++)
            for (x = x0; x <= x1; x++) {
                const float dx = ((float) x - cx) * uxf, dy = ((float) y - cy) * uyf,
                            dz = ((float) z - cz) * uzf;
                const float sq = dx * dx + dy * dy + dz * dz;
                float vk[3], vb[3], g[3], gr[3], bary[3], kk, wt, mag, fr[3];
                int face = -1, f, ddx, ddy, ddz;
                if (sq > rad * rad) continue;
                /* Rt * v with Rt = R^T (sift.c:1462-1480) */
                vk[0] = R[0] * dx + R[3] * dy + R[6] * dz;
                vk[1] = R[1] * dx + R[4] * dy + R[7] * dz;
                vk[2] = R[2] * dx + R[5] * dy + R[8] * dz;
                for (i = 0; i < 3; i++) vb[i] = (vk[i] + half_w) * bin_fctr;
                if (vb[0] < 0 || vb[1] < 0 || vb[2] < 0 || vb[0] >= 4.0f || vb[1] >= 4.0f || vb[2] >= 4.0f)
                    continue;
                grad_iso(L, x, y, z, g);
                wt = expf(-0.5f * sq / (sigma * sigma));
                g[0] = g[0] * wt; g[1] = g[1] * wt; g[2] = g[2] * wt;
                gr[0] = R[0] * g[0] + R[3] * g[1] + R[6] * g[2];
                gr[1] = R[1] * g[0] + R[4] * g[1] + R[7] * g[2];
                gr[2] = R[2] * g[0] + R[5] * g[1] + R[8] * g[2];
                if (gr[0] * gr[0] + gr[1] * gr[1] + gr[2] * gr[2] < bary_eps) continue;
                for (f = 0; f < 20; f++) {
                    if (cart2bary(gr, &g_mesh[f], bary, &kk)) continue;
                    if (bary[0] < -bary_eps || bary[1] < -bary_eps || bary[2] < -bary_eps || kk < 0) continue;
                    face = f;
                    break;
                }
                if (face < 0) continue;
                mag = sqrtf(gr[0] * gr[0] + gr[1] * gr[1] + gr[2] * gr[2]);
                for (i = 0; i < 3; i++) fr[i] = vb[i] - floorf(vb[i]);
                for (ddx = 0; ddx < 2; ddx++)
                    for (ddy = 0; ddy < 2; ddy++)
                        for (ddz = 0; ddz < 2; ddz++) {
                            const int hx = (int) vb[0] + ddx, hy = (int) vb[1] + ddy, hz = (int) vb[2] + ddz;
                            float w3, *h;
                            if (hx < 0 || hx >= 4 || hy < 0 || hy >= 4 || hz < 0 || hz >= 4) continue;
                            w3 = (ddx == 0 ? 1.0f - fr[0] : fr[0]) * (ddy == 0 ? 1.0f - fr[1] : fr[1]) *
                                 (ddz == 0 ? 1.0f - fr[2] : fr[2]);
                            h = hist + (hx + hy * 4 + hz * 16) * 12;
                            /* weights of tri.v land in tri.idx bins (the v0/v1 swap, sift.c:237-247) */
                            h[g_mesh[face].idx[0]] += mag * w3 * bary[0];
                            h[g_mesh[face].idx[1]] += mag * w3 * bary[1];
                            h[g_mesh[face].idx[2]] += mag * w3 * bary[2];
                        }
            }
    for (pass = 0; pass < 2; pass++) {
        double norm = 0.0;
        float norm_inv;
        for (i = 0; i < ORC_DESC_NUMEL; i++) norm += (double) hist[i] * hist[i];
        norm = sqrt(norm) + DBL_EPSILON;
        norm_inv = 1.0f / norm;
        for (i = 0; i < ORC_DESC_NUMEL; i++) {
            hist[i] *= norm_inv;
            if (pass == 0) hist[i] = hist[i] < trunc ? hist[i] : trunc;
        }
    }
}

int orc_describe(const orc_detector *d, const orc_key *keys, int num, float *desc) {
    int i;
    if (!d->levels || num < 1) return -1;
    if (!g_mesh_ready) build_mesh();
    for (i = 0; i < num; i++)
        if (keys[i].o < 0 || keys[i].o >= d->num_octaves || keys[i].s < -1 || keys[i].s > d->num_levels - 2)
            return -1;
#pragma omp parallel for schedule(dynamic, 1)
    for (i = 0; i < num; i++)
        describe_one(LV(d, keys[i].o, keys[i].s), &keys[i], desc + (size_t) i * ORC_DESC_NUMEL);
    return 0;
}
