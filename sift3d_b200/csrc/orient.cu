// orient.cu -- structure-tensor orientation assignment, one CTA per candidate, with an
// in-kernel symmetric 3x3 eigensolver in double (replaces the per-keypoint LAPACK dsyevd_
// call of the reference, imutil.c:984-1067).
//
// Reference semantics (sift3d/sift.c):
//   assign_orientations        :1109-1167  sigma = ori_sig_fctr * sd, per candidate
//   assign_eig_ori             :926-1085   window loop, tensor (double), window gradient
//                                          (float), rejection tests, R construction
//   IM_LOOP_SPHERE_START       :86-109     window bounds and physical-unit displacement
//   IM_GET_GRAD_ISO            :140-145    central differences / units
//   assign_orientation_thresh  :1091-1102  reject if corner score < corner_thresh
#include <string.h>
#include <float.h>

#include "common.cuh"

namespace s3d {

constexpr double kMaxEigRatio = 0.90;     // sift.c:38
constexpr double kOriGradThresh = 1E-10;  // sift.c:39
constexpr double kOriSigFctr = 1.5;       // sift.c:41
constexpr double kOriRadFctr = 3.0;       // sift.c:42

// Cyclic Jacobi eigen-decomposition of a symmetric 3x3 in double. Eigenvalues ascending
// in w (as dsyevd returns them), eigenvectors in the columns of V (as Q after the
// transpose at imutil.c:1052).
__device__ void eig3_sym(const double Ain[3][3], double V[3][3], double w[3]) {
    double A[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
            A[i][j] = Ain[i][j];
            V[i][j] = i == j ? 1.0 : 0.0;
        }
    for (int sweep = 0; sweep < 64; sweep++) {
        const double off = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
        if (off == 0.0) break;
#pragma unroll
        for (int pq = 0; pq < 3; pq++) {
            const int p = pq == 2 ? 1 : 0;
            const int q = pq == 0 ? 1 : 2;
            const double apq = A[p][q];
            if (apq == 0.0) continue;
            const double app = A[p][p], aqq = A[q][q];
            // Negligible rotation: |apq| below the resolution of both diagonal entries.
            if (fabs(apq) < DBL_MIN ||
                (fabs(app) + fabs(apq) == fabs(app) && fabs(aqq) + fabs(apq) == fabs(aqq))) {
                A[p][q] = A[q][p] = 0.0;
                continue;
            }
            const double theta = (aqq - app) / (2.0 * apq);
            const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            const double c = 1.0 / sqrt(t * t + 1.0);
            const double s = t * c;
            const double tau = s / (1.0 + c);
            A[p][p] = app - t * apq;
            A[q][q] = aqq + t * apq;
            A[p][q] = A[q][p] = 0.0;
            const int r = 3 - p - q;  // the remaining index
            const double arp = A[r][p], arq = A[r][q];
            A[r][p] = A[p][r] = arp - s * (arq + tau * arp);
            A[r][q] = A[q][r] = arq + s * (arp - tau * arq);
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const double vkp = V[k][p], vkq = V[k][q];
                V[k][p] = vkp - s * (vkq + tau * vkp);
                V[k][q] = vkq + s * (vkp - tau * vkq);
            }
        }
    }
    w[0] = A[0][0];
    w[1] = A[1][1];
    w[2] = A[2][2];
    // sort ascending (3-element network), permuting the columns of V alongside
#define S3D_SWAP_COLS(a, b)                                   \
    if (w[a] > w[b]) {                                        \
        double tw = w[a]; w[a] = w[b]; w[b] = tw;             \
        for (int k = 0; k < 3; k++) {                         \
            double tv = V[k][a]; V[k][a] = V[k][b]; V[k][b] = tv; \
        }                                                     \
    }
    S3D_SWAP_COLS(0, 1)
    S3D_SWAP_COLS(1, 2)
    S3D_SWAP_COLS(0, 1)
#undef S3D_SWAP_COLS
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}


constexpr int kOrientBatch = 16;  // windows per CTA between two finishing steps (<= 32)
template <int kOrientThreads>
__global__ void __launch_bounds__(kOrientThreads)
orient_all_kernel(const Cand *__restrict__ sorted, const OrientPlan *__restrict__ plan,
                  const int *__restrict__ counts, int *__restrict__ work, double corner_thresh,
                  sift3d_cuda_keypoint *__restrict__ out, int o_begin, int o_end,
                  sift3d_cuda_candidate_flag *__restrict__ flags) {
    __shared__ int s_first[SIFT3D_CUDA_MAX_OCTAVES + 1];  // rank of each octave's first candidate
    __shared__ int s_ci;
    __shared__ double red[kOrientThreads / 32][9];
    __shared__ double sums[kOrientBatch][9];
    const int noct = plan->noct;
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int o = 0; o < noct; o++) {
            s_first[o] = acc;
            // only this launch's octaves are counted: the others may still be scanning
            if (o >= o_begin && o < o_end) acc += min(counts[o], plan->cap[o]);
        }
        s_first[noct] = acc;
    }
    __syncthreads();
    // candidates of octaves [o_begin, o_end): ranks [s_first[o_begin], s_first[o_end])
    const int total = s_first[min(o_end, noct)];
    const int first = s_first[min(o_begin, noct)];
    // Candidates are taken in batches: the CTA sums the windows of a batch one after the other,
    // then threads 0 .. nb-1 finish one candidate each (eigen-decomposition in double, tests,
    // R) in lockstep. Finished by a single thread after every window, the serial part (~25 us
    // of double divisions and square roots) was three quarters of the kernel.
    const int batch = max(1, min(kOrientBatch, (total - first) / (int) gridDim.x));
  for (;;) {
    // dynamic hand-out: windows differ 4x in size between keypoint levels
    __syncthreads();
    if (threadIdx.x == 0) s_ci = first + atomicAdd(work, batch);
    __syncthreads();
    const int ci0 = s_ci;
    if (ci0 >= total) break;
    const int nb = min(batch, total - ci0);
   for (int b = 0; b < nb; b++) {
    const int ci = ci0 + b;
    int oi = 0;
    while (ci >= s_first[oi + 1]) oi++;
    const OctaveInfo &info = plan->oct[oi];
    const Cand cd = sorted[plan->base[oi] + (ci - s_first[oi])];
    const int s = (int) (cd.key >> 60);
    const int kz = (int) ((cd.key >> 40) & 0xFFFFF);
    const int ky = (int) ((cd.key >> 20) & 0xFFFFF);
    const int kx = (int) (cd.key & 0xFFFFF);
    const float *__restrict__ img = info.lvl[s];
    const int nx = info.nx, ny = info.ny, nz = info.nz;
    const float uxf = info.ux, uyf = info.uy, uzf = info.uz;
    const double sd = info.sd[s];
    const double sigma = kOriSigFctr * sd;      // sift.c:1125
    const double rad = sigma * kOriRadFctr;     // sift.c:936
    const float cx = (float) kx, cy = (float) ky, cz = (float) kz;  // sift.c:1124

    // Window bounds (sift.c:91-99): the double expression is narrowed to float by
    // floorf/ceilf, compared as float against the int bound, then truncated to int.
    const int x0 = (int) fmaxf(floorf((float) ((double) cx - rad / (double) uxf)), 1.0f);
    const int x1 = (int) fminf(ceilf((float) ((double) cx + rad / (double) uxf)), (float) (nx - 2));
    const int y0 = (int) fmaxf(floorf((float) ((double) cy - rad / (double) uyf)), 1.0f);
    const int y1 = (int) fminf(ceilf((float) ((double) cy + rad / (double) uyf)), (float) (ny - 2));
    const int z0 = (int) fmaxf(floorf((float) ((double) cz - rad / (double) uzf)), 1.0f);
    const int z1 = (int) fminf(ceilf((float) ((double) cz + rad / (double) uzf)), (float) (nz - 2));
    const int wx = max(x1 - x0 + 1, 0), wy = max(y1 - y0 + 1, 0), wz = max(z1 - z0 + 1, 0);
    const int total = wx * wy * wz;
    const double rad2 = rad * rad;
    const double sig2 = sigma * sigma;
    const double nhs = -0.5 / sig2;  // one division per candidate instead of one per voxel
    const float ixu = __fdiv_rn(1.0f, uxf), iyu = __fdiv_rn(1.0f, uyf), izu = __fdiv_rn(1.0f, uzf);
    const size_t ys = (size_t) nx, zs = (size_t) nx * ny;

    double a00 = 0, a01 = 0, a02 = 0, a11 = 0, a12 = 0, a22 = 0;
    double gx = 0, gy = 0, gz = 0;  // window gradient, summed in double, narrowed at the end
    // One (y, z) row of the window per thread at a time; within a row only the chord that
    // can pass the sphere test is walked (the exact test stays per voxel).
    (void) total;
    const OrientEntry *__restrict__ tab = plan->tab[oi][s];
    if (tab != nullptr) {
        // Tabulated window: the sphere around integer coordinates is the same set of offsets,
        // with the same weights, for every candidate of this level (build_orient_tables_kernel
        // evaluates the reference's test and weight once per offset). Flat list: every lane has
        // work, no distance or exp arithmetic per voxel.
        const int ne = plan->ntab[oi * SIFT3D_CUDA_MAX_FILTERS + s];
        // Branch-free body, four entries in flight per thread: an entry outside the window
        // clamps of IM_LOOP_SPHERE_START (sift.c:91-99), or past the end of the table, reads
        // the candidate's own voxel with weight 0 (adds exactly +0), so that the compiler can
        // issue the 4 x 6 loads of an unrolled step together -- the loop is bound by the
        // latency of table entry -> voxel loads, not by arithmetic.
        const float *__restrict__ q0 = img + (size_t) kz * zs + (size_t) ky * ys + (size_t) kx;
        const ptrdiff_t pys = (ptrdiff_t) ys, pzs = (ptrdiff_t) zs;
        const int nsteps = (ne + kOrientThreads - 1) / kOrientThreads;
#pragma unroll 4
        for (int k = 0; k < nsteps; k++) {
            const int e = k * kOrientThreads + threadIdx.x;
            int4 E = make_int4(0, 0, 0, 0);
            if (e < ne) E = __ldg(reinterpret_cast<const int4 *>(tab + e));
            const int x = kx + E.x, y = ky + E.y, z = kz + E.z;
            const bool in = x >= 1 && x <= nx - 2 && y >= 1 && y <= ny - 2 && z >= 1 && z <= nz - 2;
            const float w = in ? __int_as_float(E.w) : 0.0f;
            const ptrdiff_t off = in ? (ptrdiff_t) E.z * pzs + (ptrdiff_t) E.y * pys + (ptrdiff_t) E.x : 0;
            const float *__restrict__ q = q0 + off;
            float vx = __fmul_rn(0.5f, __fsub_rn(q[1], q[-1]));
            float vy = __fmul_rn(0.5f, __fsub_rn(q[pys], q[-pys]));
            float vz = __fmul_rn(0.5f, __fsub_rn(q[pzs], q[-pzs]));
            vx = __fmul_rn(vx, ixu);
            vy = __fmul_rn(vy, iyu);
            vz = __fmul_rn(vz, izu);
            const double wd = (double) w;
            const double dvx = (double) vx, dvy = (double) vy, dvz = (double) vz;
            const double wx = dvx * wd, wy = dvy * wd, wz = dvz * wd;
            a00 = fma(dvx, wx, a00);
            a01 = fma(dvx, wy, a01);
            a02 = fma(dvx, wz, a02);
            a11 = fma(dvy, wy, a11);
            a12 = fma(dvy, wz, a12);
            a22 = fma(dvz, wz, a22);
            gx += (double) __fmul_rn(vx, w);
            gy += (double) __fmul_rn(vy, w);
            gz += (double) __fmul_rn(vz, w);
        }
    }
    const float rad2f_hi = (float) rad2 * 1.0001f;
    const int nrows = (tab == nullptr && wx > 0) ? wy * wz : 0;
    for (int r = threadIdx.x; r < nrows; r += kOrientThreads) {
        const int y = y0 + r % wy, z = z0 + r / wy;
        const float dy = __fmul_rn(__fsub_rn((float) y, cy), uyf);
        const float dz = __fmul_rn(__fsub_rn((float) z, cz), uzf);
        const float dy2 = __fmul_rn(dy, dy), dz2 = __fmul_rn(dz, dz);
        const float rem = rad2f_hi - dy2 - dz2;
        if (rem < 0.0f) continue;
        const float half = sqrtf(rem) * ixu + 0.01f;
        const int xa = max(x0, (int) ceilf(cx - half)), xb = min(x1, (int) floorf(cx + half));
        const float *__restrict__ row = img + (size_t) z * zs + (size_t) y * ys;
        for (int x = xa; x <= xb; x++) {
            const float dx = __fmul_rn(__fsub_rn((float) x, cx), uxf);
            const float sq = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), dy2), dz2);
            if ((double) sq > rad2) continue;
            // weight = expf(-0.5 * sq_dist / (sigma * sigma)): the argument is evaluated in
            // double and narrowed to float (sift.c:972); expf itself in fp32 (<= 2 ulp)
            const float arg = (float) ((double) sq * nhs);  // -0.5 * sq / sigma^2 in double
            const float w = expf(arg);
            float vx = __fmul_rn(0.5f, __fsub_rn(row[x + 1], row[x - 1]));
            float vy = __fmul_rn(0.5f, __fsub_rn(row[x + ys], row[(ptrdiff_t) x - (ptrdiff_t) ys]));
            float vz = __fmul_rn(0.5f, __fsub_rn(row[x + zs], row[(ptrdiff_t) x - (ptrdiff_t) zs]));
            vx = __fmul_rn(vx, ixu);
            vy = __fmul_rn(vy, iyu);
            vz = __fmul_rn(vz, izu);
            const double wd = (double) w;
            const double dvx = (double) vx, dvy = (double) vy, dvz = (double) vz;
            // structure tensor (sift.c:978-986) with the weight folded in once per row
            const double wx = dvx * wd, wy = dvy * wd, wz = dvz * wd;
            a00 = fma(dvx, wx, a00);
            a01 = fma(dvx, wy, a01);
            a02 = fma(dvx, wz, a02);
            a11 = fma(dvy, wy, a11);
            a12 = fma(dvy, wz, a12);
            a22 = fma(dvz, wz, a22);
            gx += (double) __fmul_rn(vx, w);
            gy += (double) __fmul_rn(vy, w);
            gz += (double) __fmul_rn(vz, w);
        }
    }

    double vals[9] = {a00, a01, a02, a11, a12, a22, gx, gy, gz};
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const double r = warp_sum(vals[k]);
        if (lane == 0) red[warp][k] = r;
    }
    __syncthreads();
    if (threadIdx.x < 9) {
        double r = 0;
        for (int wv = 0; wv < kOrientThreads / 32; wv++) r += red[wv][threadIdx.x];
        sums[b][threadIdx.x] = r;
    }
    __syncthreads();  // red is reused by the next window
   }  // windows of the batch
    if (threadIdx.x >= nb) continue;
    // ---- one candidate per thread from here on
    const int ci = ci0 + threadIdx.x;
    int oi = 0;
    while (ci >= s_first[oi + 1]) oi++;
    const OctaveInfo &info = plan->oct[oi];
    const Cand cd = sorted[plan->base[oi] + (ci - s_first[oi])];
    const int s = (int) (cd.key >> 60);
    const int kz = (int) ((cd.key >> 40) & 0xFFFFF);
    const int ky = (int) ((cd.key >> 20) & 0xFFFFF);
    const int kx = (int) (cd.key & 0xFFFFF);
    const double sd = info.sd[s];
    double vals[9];
#pragma unroll
    for (int k = 0; k < 9; k++) vals[k] = sums[threadIdx.x][k];

    sift3d_cuda_keypoint kp;
    kp.o = info.o;
    kp.s = s;
    kp.x = (double) kx;
    kp.y = (double) ky;
    kp.z = (double) kz;
    kp.sd = sd;
    kp.strength = cd.strength;
    for (int k = 0; k < 9; k++) kp.R[k] = 0.0f;
    kp.valid = 0;

    const float wgx = (float) vals[6], wgy = (float) vals[7], wgz = (float) vals[8];
    const float gnorm2 = __fadd_rn(__fadd_rn(__fmul_rn(wgx, wgx), __fmul_rn(wgy, wgy)), __fmul_rn(wgz, wgz));
    bool ok = !(gnorm2 < (float) kOriGradThresh);  // sift.c:997
    if (ok) {
        double A[3][3] = {{vals[0], vals[1], vals[2]}, {vals[1], vals[3], vals[4]}, {vals[2], vals[4], vals[5]}};
        double V[3][3], L[3];
        eig3_sym(A, V, L);
        // eigenvalue-ratio stability test (sift.c:1011-1015)
        if (fabs(L[0] / L[1]) > kMaxEigRatio || fabs(L[1] / L[2]) > kMaxEigRatio) ok = false;
        if (ok) {
            double corner = DBL_MAX;
            float v0[3], v1[3];
            const float gnorm = sqrtf(gnorm2);
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const int e = 2 - i;  // descending eigenvalue order (sift.c:1021)
                float vr0 = (float) V[0][e], vr1 = (float) V[1][e], vr2 = (float) V[2][e];
                const double d = (double) __fadd_rn(__fadd_rn(__fmul_rn(wgx, vr0), __fmul_rn(wgy, vr1)), __fmul_rn(wgz, vr2));
                const float vnorm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(vr0, vr0), __fmul_rn(vr1, vr1)), __fmul_rn(vr2, vr2)));
                const double cosang = d / (double) __fmul_rn(vnorm, gnorm);
                corner = fmin(corner, fabs(cosang));
                const float sgn = d > 0.0 ? 1.0f : -1.0f;
                vr0 = __fmul_rn(vr0, sgn);
                vr1 = __fmul_rn(vr1, sgn);
                vr2 = __fmul_rn(vr2, sgn);
                kp.R[0 * 3 + i] = vr0;
                kp.R[1 * 3 + i] = vr1;
                kp.R[2 * 3 + i] = vr2;
                if (i == 0) { v0[0] = vr0; v0[1] = vr1; v0[2] = vr2; }
                else        { v1[0] = vr0; v1[1] = vr1; v1[2] = vr2; }
            }
            // third axis = v0 x v1 (sift.c:1054-1059)
            kp.R[0 * 3 + 2] = __fsub_rn(__fmul_rn(v0[1], v1[2]), __fmul_rn(v0[2], v1[1]));
            kp.R[1 * 3 + 2] = __fsub_rn(__fmul_rn(v0[2], v1[0]), __fmul_rn(v0[0], v1[2]));
            kp.R[2 * 3 + 2] = __fsub_rn(__fmul_rn(v0[0], v1[1]), __fmul_rn(v0[1], v1[0]));
            if (corner < corner_thresh) ok = false;  // sift.c:1100-1101
        }
    }
    kp.valid = ok ? 1 : 0;
    // results stay in the octave's own region (slot = rank within the octave): octaves are
    // oriented independently, as each finishes, without knowing the counts of the others
    out[plan->base[oi] + (ci - s_first[oi])] = kp;
    if (flags != nullptr) {
        sift3d_cuda_candidate_flag f;
        f.strength = kp.strength;
        f.valid = kp.valid;
        flags[plan->base[oi] + (ci - s_first[oi])] = f;
    }
  }  // candidates
}

// One CTA per (octave, keypoint level): offsets (z, y, x ascending, the reference's loop order)
// of the voxels inside the sphere, with their weights -- the same expressions as the per-voxel
// path of orient_all_kernel, evaluated for a candidate at the origin ((float) x - cx is the
// exact integer offset).
struct OrientTableJob {
    OrientEntry *out;
    int *count;
    int rx, ry, rz;  // half extents of the box that is scanned
    float ux, uy, uz;
    double sd;
};
struct OrientTableJobs {
    OrientTableJob job[64];
};
constexpr int kTabThreads = 256;
__global__ void __launch_bounds__(kTabThreads)
build_orient_tables_kernel(const OrientTableJobs J) {
    const OrientTableJob &T = J.job[blockIdx.x];
    __shared__ int wcount[kTabThreads / 32];
    __shared__ int base_sh;
    const double sigma = kOriSigFctr * T.sd;
    const double rad = sigma * kOriRadFctr;
    const double rad2 = rad * rad;
    const double nhs = -0.5 / (sigma * sigma);
    const int bx = 2 * T.rx + 1, by = 2 * T.ry + 1, bz = 2 * T.rz + 1;
    const int total = bx * by * bz;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) base_sh = 0;
    __syncthreads();
    for (int i0 = 0; i0 < total; i0 += kTabThreads) {
        const int i = i0 + threadIdx.x;
        bool in = false;
        int dxi = 0, dyi = 0, dzi = 0;
        float w = 0.0f;
        if (i < total) {
            dxi = i % bx - T.rx;
            dyi = (i / bx) % by - T.ry;
            dzi = i / (bx * by) - T.rz;
            const float dx = __fmul_rn((float) dxi, T.ux);
            const float dy = __fmul_rn((float) dyi, T.uy);
            const float dz = __fmul_rn((float) dzi, T.uz);
            const float sq = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            in = !((double) sq > rad2);
            w = expf((float) ((double) sq * nhs));
        }
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (lane == 0) wcount[warp] = __popc(m);
        __syncthreads();
        int before = base_sh;
        for (int k = 0; k < warp; k++) before += wcount[k];
        if (in) {
            OrientEntry e;
            e.dx = dxi; e.dy = dyi; e.dz = dzi; e.w = w;
            T.out[before + __popc(m & ((1u << lane) - 1u))] = e;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int k = 0; k < kTabThreads / 32; k++) tot += wcount[k];
            base_sh += tot;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *T.count = base_sh;
}

static void orient_box(const OctaveInfo &info, int s, int r[3]) {
    const double rad = kOriSigFctr * info.sd[s] * kOriRadFctr;
    r[0] = (int) ceil(rad / (double) info.ux) + 1;
    r[1] = (int) ceil(rad / (double) info.uy) + 1;
    r[2] = (int) ceil(rad / (double) info.uz) + 1;
}

size_t orient_table_capacity(const OrientPlan &h, int kp,
                             size_t off[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS]) {
    size_t total = 0;
    if (h.noct * kp > 64) return 0;  // OrientTableJobs holds 64 tables
    for (int o = 0; o < h.noct; o++)
        for (int s = 0; s < kp; s++) {
            int r[3];
            orient_box(h.oct[o], s, r);
            const double vol = (2.0 * r[0] + 1) * (2.0 * r[1] + 1) * (2.0 * r[2] + 1);
            if (vol > 4.0e6) return 0;  // absurd units / scales: keep the per-voxel path
            off[o][s] = total;
            total += ((size_t) vol + 3) & ~(size_t) 3;  // 16-byte aligned tables
        }
    return total;
}

int launch_orient_tables(cudaStream_t st, OrientPlan &h, int kp, OrientEntry *d_entries, int *d_counts,
                         uint64_t *launches) {
    size_t off[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS];
    if (orient_table_capacity(h, kp, off) == 0) return 1;
    OrientTableJobs J;
    int nj = 0;
    for (int o = 0; o < h.noct; o++)
        for (int s = 0; s < kp; s++) {
            OrientTableJob &T = J.job[nj++];
            int r[3];
            orient_box(h.oct[o], s, r);
            T.out = d_entries + off[o][s];
            T.count = d_counts + o * SIFT3D_CUDA_MAX_FILTERS + s;
            T.rx = r[0]; T.ry = r[1]; T.rz = r[2];
            T.ux = h.oct[o].ux; T.uy = h.oct[o].uy; T.uz = h.oct[o].uz;
            T.sd = h.oct[o].sd[s];
            h.tab[o][s] = T.out;
        }
    h.ntab = d_counts;
    build_orient_tables_kernel<<<nj, kTabThreads, 0, st>>>(J);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

int launch_orient_all(cudaStream_t st, const Cand *d_sorted, const OrientPlan *d_plan,
                      const int *d_counts, int *d_work, double corner_thresh,
                      sift3d_cuda_keypoint *d_out, uint64_t *launches, int o_begin, int o_end,
                      sift3d_cuda_candidate_flag *d_flags, bool shared_device) {
    if (o_end < 0) o_end = SIFT3D_CUDA_MAX_OCTAVES;
    if (o_end <= o_begin) return 0;
    // persistent grid: 148 SMs x the CTAs that fit (94 registers x 128 threads -> 5 per SM)
    // Octave 0 on the main stream: persistent grid (148 SMs x the CTAs that fit). Other octaves
    // run on side streams while octave 0's one-wave blurs still need whole SMs: many short
    // CTAs (about one candidate each) so that an SM is never held for long.
    // The small octaves are the tail of detect and hold few candidates: there the time of ONE
    // candidate counts (6.3k window voxels over 128 threads = 50 dependent load steps, 40 us),
    // so their CTAs are four times as wide.
    static const int wide_env = getenv("SIFT3D_CUDA_ORIENT_WIDE") ? atoi(getenv("SIFT3D_CUDA_ORIENT_WIDE")) : 0;
    // The kernel is latency-bound (table entry -> voxel loads) and its persistent grid reserves
    // the registers of the SMs it sits on: 5 CTAs per SM (8 short ones for the other octaves) is
    // the fastest for a detector that has the device to itself (2.30 against 2.33 ms per 256^3
    // volume), 2 per SM leaves room for the kernels of the other pipelines of the device (four
    // pipelines: 9.04 against 8.77 Gvox/s). $SIFT3D_CUDA_ORIENT_CTAS="a,b" forces both values.
    static const char *ce = getenv("SIFT3D_CUDA_ORIENT_CTAS");
    const int cta0 = ce ? atoi(ce) : (shared_device ? 2 : 5);
    const int cta1 = ce && strchr(ce, ',') ? atoi(strchr(ce, ',') + 1) : (shared_device ? 2 : 8);
    if (o_begin == 0 || !wide_env)
        orient_all_kernel<128><<<o_begin == 0 ? kNumSMs * cta0 : kNumSMs * cta1, 128, 0, st>>>(
            d_sorted, d_plan, d_counts, d_work, corner_thresh, d_out, o_begin, o_end, d_flags);
    else
        orient_all_kernel<512><<<kNumSMs * 2, 512, 0, st>>>(d_sorted, d_plan, d_counts, d_work, corner_thresh,
                                                           d_out, o_begin, o_end, d_flags);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

}  // namespace s3d
