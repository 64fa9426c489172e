/* synth.c -- deterministic synthetic volumes for the parity tests and bench.py.
 *
 * The reference ships no sample data, so the workloads of
 * BASELINE.json `configs` are generated here: additive Gaussian blobs
 * ("C1" 128^3), blobs x smooth bias + noise ("C2" MRI-like 256^3), and the
 * anisotropic CT-shaped volume ("C3"). The same buffer feeds the CUDA path and
 * the CPU oracle, so only determinism matters, not the exact recipe.
 *
 * PRNG: xorshift64 (13, 7, 17), u = (s >> 11) * 2^-53.
 * Per blob, in this draw order: cx = u*nx, cy = u*ny, cz = u*nz,
 * sigma = 1.5 + 4.5u, sign = (u < 0.5 ? -1 : +1), amp = 0.3 + 0.7u.
 * Support r = (int)(3 sigma) + 1 around (int)c, clipped to the volume;
 * voxel += (float)(((sign*amp * ez[z]) * ey[y]) * ex[x]),
 * e?[i] = exp(-(i - c?)^2 / (2 sigma^2)) in double. Blobs are applied in draw
 * order at every voxel (threads split the volume by z, not the blob list), so
 * the result does not depend on the thread count.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SYNTH_EXPORT __attribute__((visibility("default")))

typedef struct {
    double cx, cy, cz, sigma, sa;
    int r;
} blob_t;

static inline double xs_next(uint64_t *s) {
    uint64_t x = *s;
    x ^= x << 13;
    x ^= x >> 7;
    x ^= x << 17;
    *s = x;
    return (double) (x >> 11) * (1.0 / 9007199254740992.0);
}

/* flags: bit0 = multiply by the smooth bias field 1 + 0.3 cos(2 pi x/nx) cos(2 pi y/ny),
 *        bit1 = add N(0, 0.02) noise (Box-Muller, drawn after all blobs, voxel order).
 * Returns 0 on success. */
SYNTH_EXPORT int sift3d_synth_volume(float *vol, int nx, int ny, int nz, int num_blobs,
                                     uint64_t seed, int flags) {
    const size_t nvox = (size_t) nx * ny * nz;
    blob_t *blobs;
    uint64_t s = seed ? seed : 88172645463325252ULL;
    int b;

    if (vol == NULL || nx < 1 || ny < 1 || nz < 1 || num_blobs < 0)
        return -1;
    blobs = (blob_t *) malloc(sizeof(blob_t) * (size_t) (num_blobs > 0 ? num_blobs : 1));
    if (blobs == NULL)
        return -1;
    for (b = 0; b < num_blobs; b++) {
        double sign, amp;
        blobs[b].cx = xs_next(&s) * nx;
        blobs[b].cy = xs_next(&s) * ny;
        blobs[b].cz = xs_next(&s) * nz;
        blobs[b].sigma = 1.5 + 4.5 * xs_next(&s);
        sign = xs_next(&s) < 0.5 ? -1.0 : 1.0;
        amp = 0.3 + 0.7 * xs_next(&s);
        blobs[b].sa = sign * amp;
        blobs[b].r = (int) (3.0 * blobs[b].sigma) + 1;
    }
    memset(vol, 0, nvox * sizeof(float));

#pragma omp parallel
    {
        double ex[64], ey[64];
        int z;
#pragma omp for schedule(static)
        for (z = 0; z < nz; z++) {
            int bb;
            for (bb = 0; bb < num_blobs; bb++) {
                const blob_t *const B = blobs + bb;
                const int r = B->r;
                const int x0 = (int) B->cx - r < 0 ? 0 : (int) B->cx - r;
                const int x1 = (int) B->cx + r > nx - 1 ? nx - 1 : (int) B->cx + r;
                const int y0 = (int) B->cy - r < 0 ? 0 : (int) B->cy - r;
                const int y1 = (int) B->cy + r > ny - 1 ? ny - 1 : (int) B->cy + r;
                const double inv = 1.0 / (2.0 * B->sigma * B->sigma);
                double ez, dz;
                int x, y;
                if (z < (int) B->cz - r || z > (int) B->cz + r)
                    continue;
                dz = (double) z - B->cz;
                ez = B->sa * exp(-dz * dz * inv);
                for (x = x0; x <= x1; x++) {
                    const double dx = (double) x - B->cx;
                    ex[x - x0] = exp(-dx * dx * inv);
                }
                for (y = y0; y <= y1; y++) {
                    const double dy = (double) y - B->cy;
                    ey[y - y0] = exp(-dy * dy * inv);
                }
                for (y = y0; y <= y1; y++) {
                    float *row = vol + ((size_t) z * ny + y) * nx;
                    const double ezy = ez * ey[y - y0];
                    for (x = x0; x <= x1; x++)
                        row[x] += (float) (ezy * ex[x - x0]);
                }
            }
        }
    }
    free(blobs);

    if (flags & 1) {
        const double two_pi = 6.283185307179586476925286766559;
        int x, y, z;
#pragma omp parallel for private(x, y) schedule(static)
        for (z = 0; z < nz; z++)
            for (y = 0; y < ny; y++) {
                const double by = cos(two_pi * y / ny);
                float *row = vol + ((size_t) z * ny + y) * nx;
                for (x = 0; x < nx; x++)
                    row[x] = (float) (row[x] * (1.0 + 0.3 * cos(two_pi * x / nx) * by));
            }
    }
    if (flags & 2) {
        const double two_pi = 6.283185307179586476925286766559;
        size_t i;
        for (i = 0; i + 1 < nvox; i += 2) {
            double u1 = xs_next(&s), u2 = xs_next(&s), m;
            if (u1 < 1e-300) u1 = 1e-300;
            m = 0.02 * sqrt(-2.0 * log(u1));
            vol[i] += (float) (m * cos(two_pi * u2));
            vol[i + 1] += (float) (m * sin(two_pi * u2));
        }
    }
    return 0;
}
