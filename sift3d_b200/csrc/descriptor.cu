// descriptor.cu -- icosahedral gradient-orientation histogram descriptor, one CTA per
// keypoint with the 4x4x4x12 histogram in shared memory.
//
// Reference semantics (sift3d/sift.c):
//   init_geometry           :148-259    icosahedron; the v[0]<->v[1] swap without idx swap
//   cart2bary               :268-327    Moller-Trumbore barycentric coordinates
//   icos_hist_bin           :1254-1291  first face (index order) hit by the gradient ray
//   SIFT3D_desc_acc_interp  :1295-1399  2x2x2 spatial x 3 barycentric weights
//   normalize_desc          :1402-1429  L2 norm in double, + DBL_EPSILON
//   extract_descrip         :1442-1536  window loop, rotate, weight, accumulate,
//                                       normalise / clamp to 0.2*128/768 / normalise
#include <cooperative_groups.h>
#include <float.h>
#include <math.h>

#include "common.cuh"

namespace s3d {

constexpr int kFaces = 20;
constexpr int kVerts = 12;
constexpr int kHistPerDim = 4;
constexpr int kDescNumel = SIFT3D_CUDA_DESC_NUMEL;
constexpr double kBaryEps = (double) FLT_EPSILON * 1E1;  // sift.c:40
// 10 * 2^-23 is a float, so for a float x the reference's (double) x < BARY_EPS is x < kBaryEpsF
constexpr float kBaryEpsF = FLT_EPSILON * 10.0f;
static_assert((double) kBaryEpsF == kBaryEps, "BARY_EPS must be exactly representable in fp32");
constexpr double kDescRadFctr = 2.0;                     // sift.c:44

// Per-face constants of cart2bary that do not depend on the query vector.
struct MeshConst {
    float e1[kFaces][3];   // v1 - v0
    float e2[kFaces][3];   // v2 - v0
    float t[kFaces][3];    // -v0
    float q[kFaces][3];    // t x e1
    float e2q[kFaces];     // e2 . q
    float ctr[kFaces][3];  // v0 + v1 + v2 (direction of the face centre; selection only)
    float minv[kFaces][12];  // rows of [v0 v1 v2]^-1, padded to 4 (fast-path barycentric solve)
    int idx[kFaces][3];    // histogram bin of each barycentric weight
    // face whose centre is nearest to a direction, by octant (bit 0: x < 0, bit 1: y < 0,
    // bit 2: z < 0) and by which of the 4 candidate centres of that octant wins
    unsigned char lut[8][4];
};

__constant__ MeshConst c_mesh;

static inline void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
static inline float dot3(const float a[3], const float b[3]) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

// Build the table the way init_geometry does (host, fp32, x86-64 without FMA so every
// operation rounds exactly as in the reference build) and upload it.
int upload_mesh(void) {
    const double gr = 1.6180339887;  // sift.c:48
    const float vert[kVerts][3] = {
        {0, 1, (float) gr},  {0, -1, (float) gr},  {0, 1, (float) -gr}, {0, -1, (float) -gr},
        {1, (float) gr, 0},  {-1, (float) gr, 0},  {1, (float) -gr, 0}, {-1, (float) -gr, 0},
        {(float) gr, 0, 1},  {(float) -gr, 0, 1},  {(float) gr, 0, -1}, {(float) -gr, 0, -1}};
    const int faces[kFaces][3] = {{0, 1, 8},  {0, 8, 4},  {0, 4, 5},   {0, 5, 9},  {0, 9, 1},
                                  {1, 6, 8},  {8, 6, 10}, {8, 10, 4},  {4, 10, 2}, {4, 2, 5},
                                  {5, 2, 11}, {5, 11, 9}, {9, 11, 7},  {9, 7, 1},  {1, 7, 6},
                                  {3, 6, 7},  {3, 7, 11}, {3, 11, 2},  {3, 2, 10}, {3, 10, 6}};
    MeshConst m;
    for (int i = 0; i < kFaces; i++) {
        volatile float v[3][3];
        for (int j = 0; j < 3; j++) {
            const int id = faces[i][j];
            m.idx[i][j] = id;
            const float x = vert[id][0], y = vert[id][1], z = vert[id][2];
            const float mag = sqrtf(x * x + y * y + z * z);
            // SIFT3D_CVEC_SCALE(v + j, 1.0f / mag) expands to v.x * 1.0f / mag (sift.c:228)
            v[j][0] = x * 1.0f / mag;
            v[j][1] = y * 1.0f / mag;
            v[j][2] = z * 1.0f / mag;
        }
        float a[3], b[3], n[3], v0[3], v1[3], v2[3];
        for (int k = 0; k < 3; k++) { v0[k] = v[0][k]; v1[k] = v[1][k]; v2[k] = v[2][k]; }
        for (int k = 0; k < 3; k++) { a[k] = v2[k] - v1[k]; b[k] = v1[k] - v0[k]; }
        cross3(a, b, n);
        if (dot3(n, v0) < 0) {  // sift.c:237-242: vertices swapped, idx left alone
            for (int k = 0; k < 3; k++) { const float tmp = v0[k]; v0[k] = v1[k]; v1[k] = tmp; }
        }
        for (int k = 0; k < 3; k++) {
            m.e1[i][k] = v1[k] - v0[k];
            m.e2[i][k] = v2[k] - v0[k];
            m.t[i][k] = v0[k] * -1.0f;
            m.ctr[i][k] = v0[k] + v1[k] + v2[k];
        }
        cross3(m.t[i], m.e1[i], m.q[i]);
        m.e2q[i] = dot3(m.e2[i], m.q[i]);
        // [v0 v1 v2]^-1 (double, then narrowed): b' = Minv g is proportional to the barycentric
        // coordinates of the point where the ray g leaves the face, b = b' / sum(b')
        {
            const double A[3][3] = {{v0[0], v1[0], v2[0]}, {v0[1], v1[1], v2[1]}, {v0[2], v1[2], v2[2]}};
            const double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) -
                               A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                               A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
            for (int r = 0; r < 3; r++) {
                for (int c = 0; c < 3; c++) {
                    const int r1 = (r + 1) % 3, r2 = (r + 2) % 3, c1 = (c + 1) % 3, c2 = (c + 2) % 3;
                    // inverse = adjugate / det; entry (r, c) is the cofactor of (c, r)
                    m.minv[i][4 * r + c] = (float) ((A[c1][r1] * A[c2][r2] - A[c1][r2] * A[c2][r1]) / det);
                }
                m.minv[i][4 * r + 3] = 0.0f;
            }
        }
    }
    // Nearest-centre lookup. By symmetry a direction in the octant with signs (sx, sy, sz)
    // is nearest to one of 4 face centres: the octant's own face (1,1,1) and the three
    // faces straddling the octant's coordinate planes, (phi,0,2phi+1), (0,2phi+1,phi),
    // (2phi+1,phi,0) (all four of equal length). The table maps them onto the mesh's face
    // numbering by matching the actual centre vectors, so it cannot disagree with `ctr`.
    {
        const float ph = (float) gr;
        const float cand[4][3] = {{1 + ph, 1 + ph, 1 + ph}, {ph, 0, 2 * ph + 1}, {0, 2 * ph + 1, ph},
                                  {2 * ph + 1, ph, 0}};
        for (int oct = 0; oct < 8; oct++)
            for (int c = 0; c < 4; c++) {
                const float v[3] = {cand[c][0] * ((oct & 1) ? -1.f : 1.f),
                                    cand[c][1] * ((oct & 2) ? -1.f : 1.f),
                                    cand[c][2] * ((oct & 4) ? -1.f : 1.f)};
                int best = 0;
                float bd = -1e30f;
                for (int f = 0; f < kFaces; f++) {
                    const float d = dot3(v, m.ctr[f]);
                    if (d > bd) { bd = d; best = f; }
                }
                m.lut[oct][c] = (unsigned char) best;
            }
    }
    S3D_CK(cudaMemcpyToSymbol(c_mesh, &m, sizeof(m)));
    return 0;
}

// cart2bary for face f (sift.c:268-297). Returns false if the reference would `continue`
// past this face in icos_hist_bin (unstable, or the ray misses the triangle).
__device__ __forceinline__ bool face_hit(int f, float gx, float gy, float gz, float &bx,
                                         float &by, float &bz) {
    const float e2x = c_mesh.e2[f][0], e2y = c_mesh.e2[f][1], e2z = c_mesh.e2[f][2];
    // p = cart x e2
    const float px = __fsub_rn(__fmul_rn(gy, e2z), __fmul_rn(gz, e2y));
    const float py = __fsub_rn(__fmul_rn(gz, e2x), __fmul_rn(gx, e2z));
    const float pz = __fsub_rn(__fmul_rn(gx, e2y), __fmul_rn(gy, e2x));
    const float det = __fadd_rn(__fadd_rn(__fmul_rn(c_mesh.e1[f][0], px), __fmul_rn(c_mesh.e1[f][1], py)),
                                __fmul_rn(c_mesh.e1[f][2], pz));
    if ((double) fabsf(det) < kBaryEps) return false;
    const float det_inv = __fdiv_rn(1.0f, det);
    const float tp = __fadd_rn(__fadd_rn(__fmul_rn(c_mesh.t[f][0], px), __fmul_rn(c_mesh.t[f][1], py)),
                               __fmul_rn(c_mesh.t[f][2], pz));
    const float cq = __fadd_rn(__fadd_rn(__fmul_rn(gx, c_mesh.q[f][0]), __fmul_rn(gy, c_mesh.q[f][1])),
                               __fmul_rn(gz, c_mesh.q[f][2]));
    by = __fmul_rn(det_inv, tp);
    bz = __fmul_rn(det_inv, cq);
    bx = __fsub_rn(__fsub_rn(1.0f, by), bz);
    const float k = __fmul_rn(c_mesh.e2q[f], det_inv);
    if ((double) bx < -kBaryEps || (double) by < -kBaryEps || (double) bz < -kBaryEps || k < 0.0f)
        return false;
    return true;
}

constexpr int kDescThreads = 256;
constexpr float kFixScale = 1073741824.0f;          // 2^30 fixed-point units per 1.0
constexpr double kFixInv = 1.0 / 1073741824.0;

// Histogram accumulation in 64-bit fixed point held as two 32-bit words, so that the
// only shared-memory atomic on the hot path is the native 32-bit integer add
// (ATOMS.ADD); fp32 shared atomics compile to CAS spin loops (ATOMS.CAST.SPIN) and
// serialise badly because neighbouring voxels hit the same bins. Integer adds commute,
// so descriptors are bit-reproducible from run to run.
__device__ __forceinline__ void fix_add(unsigned *lo, unsigned *hi, int i, float v) {
    const long long q = __float2ll_rn(v * kFixScale);
    const unsigned ql = (unsigned) q;
    unsigned qh = (unsigned) ((unsigned long long) q >> 32);
    const unsigned old = atomicAdd(lo + i, ql);
    qh += (unsigned) (old + ql < old);  // carry out of the low word
    if (qh) atomicAdd(hi + i, qh);
}

__global__ void __launch_bounds__(kDescThreads)
descriptor_rows_kernel(const DescJob *__restrict__ jobs, int n, float *__restrict__ desc_out) {
    __shared__ unsigned hlo[kDescNumel];
    __shared__ unsigned hhi[kDescNumel];
    __shared__ float hist[kDescNumel];
    __shared__ double red[kDescThreads / 32];
    __shared__ double norm_sh;
    const int ki = blockIdx.x;
    if (ki >= n) return;
    const DescJob J = jobs[ki];
    for (int i = threadIdx.x; i < kDescNumel; i += kDescThreads) {
        hlo[i] = 0u;
        hhi[i] = 0u;
    }
    __syncthreads();

    const float *__restrict__ img = J.img;
    const int nx = J.nx, ny = J.ny, nz = J.nz;
    const float uxf = J.ux, uyf = J.uy, uzf = J.uz;
    const float cx = J.cx, cy = J.cy, cz = J.cz;
    // sift.c:1453-1458
    const float sigma = J.sigma;
    const float rad = (float) (kDescRadFctr * (double) sigma);
    const float half_w = (float) ((double) rad / sqrt(2.0));
    const float desc_w = __fmul_rn(2.0f, half_w);
    const float hist_w = __fdiv_rn(desc_w, (float) kHistPerDim);
    const float bin_fctr = __fdiv_rn(1.0f, hist_w);
    const float rad2 = __fmul_rn(rad, rad);
    const float sig2 = __fmul_rn(sigma, sigma);
    const float ixu = __fdiv_rn(1.0f, uxf), iyu = __fdiv_rn(1.0f, uyf), izu = __fdiv_rn(1.0f, uzf);

    // Window (sift.c:91-99), all-float here because win_radius is a float.
    const int x0 = (int) fmaxf(floorf(__fsub_rn(cx, __fdiv_rn(rad, uxf))), 1.0f);
    const int x1 = (int) fminf(ceilf(__fadd_rn(cx, __fdiv_rn(rad, uxf))), (float) (nx - 2));
    const int y0 = (int) fmaxf(floorf(__fsub_rn(cy, __fdiv_rn(rad, uyf))), 1.0f);
    const int y1 = (int) fminf(ceilf(__fadd_rn(cy, __fdiv_rn(rad, uyf))), (float) (ny - 2));
    const int z0 = (int) fmaxf(floorf(__fsub_rn(cz, __fdiv_rn(rad, uzf))), 1.0f);
    const int z1 = (int) fminf(ceilf(__fadd_rn(cz, __fdiv_rn(rad, uzf))), (float) (nz - 2));
    const int wy = max(y1 - y0 + 1, 0), wz = max(z1 - z0 + 1, 0);
    const int nrows = (x1 >= x0) ? wy * wz : 0;
    const size_t ys = (size_t) nx, zs = (size_t) nx * ny;

    // Rt = R^T (sift.c:1462-1464): Rt[i][j] = R[j][i]
    const float r00 = J.R[0], r01 = J.R[1], r02 = J.R[2];
    const float r10 = J.R[3], r11 = J.R[4], r12 = J.R[5];
    const float r20 = J.R[6], r21 = J.R[7], r22 = J.R[8];

    // One (y, z) row of the window per thread at a time; the thread walks the row in x.
    // Consecutive voxels mostly fall into the same (spatial cell, icosahedron face), so
    // their 8 x 3 weighted contributions are summed in registers and flushed to shared
    // memory only when that key changes.
    float acc[8][3];
    int cur_key = -1;
    int ci0 = 0, ci1 = 0, ci2 = 0, cbx = 0, cby = 0, cbz = 0;

#define S3D_FLUSH()                                                                        \
    if (cur_key >= 0) {                                                                    \
        _Pragma("unroll") for (int c = 0; c < 8; c++) {                                    \
            const int hx = cbx + (c >> 2), hy = cby + ((c >> 1) & 1), hz = cbz + (c & 1);  \
            if (hx < kHistPerDim && hy < kHistPerDim && hz < kHistPerDim) {                \
                const int hb = (hx + hy * kHistPerDim + hz * kHistPerDim * kHistPerDim) * kVerts; \
                if (acc[c][0] != 0.0f) fix_add(hlo, hhi, hb + ci0, acc[c][0]);             \
                if (acc[c][1] != 0.0f) fix_add(hlo, hhi, hb + ci1, acc[c][1]);             \
                if (acc[c][2] != 0.0f) fix_add(hlo, hhi, hb + ci2, acc[c][2]);             \
            }                                                                              \
        }                                                                                  \
    }

    for (int r = threadIdx.x; r < nrows; r += kDescThreads) {
        const int y = y0 + r % wy, z = z0 + r / wy;
        const float dy = __fmul_rn(__fsub_rn((float) y, cy), uyf);
        const float dz = __fmul_rn(__fsub_rn((float) z, cz), uzf);
        // conservative chord of the sphere along this row; the exact test stays per voxel
        const float rem = rad2 * 1.0001f - dy * dy - dz * dz;
        if (rem < 0.0f) continue;
        const float half = sqrtf(rem) * ixu + 1.0f;
        const int xa = max(x0, (int) floorf(cx - half));
        const int xb = min(x1, (int) ceilf(cx + half));
        const float *__restrict__ row = img + (size_t) z * zs + (size_t) y * ys;
        for (int x = xa; x <= xb; x++) {
            const float dx = __fmul_rn(__fsub_rn((float) x, cx), uxf);
            const float sq = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            if (sq > rad2) continue;
            // vkp = Rt * vim (SIFT3D_MUL_MAT_RM_CVEC, immacros.h:326-342)
            const float kx = __fadd_rn(__fadd_rn(__fmul_rn(r00, dx), __fmul_rn(r10, dy)), __fmul_rn(r20, dz));
            const float ky = __fadd_rn(__fadd_rn(__fmul_rn(r01, dx), __fmul_rn(r11, dy)), __fmul_rn(r21, dz));
            const float kz = __fadd_rn(__fadd_rn(__fmul_rn(r02, dx), __fmul_rn(r12, dy)), __fmul_rn(r22, dz));
            const float vbx = __fmul_rn(__fadd_rn(kx, half_w), bin_fctr);
            const float vby = __fmul_rn(__fadd_rn(ky, half_w), bin_fctr);
            const float vbz = __fmul_rn(__fadd_rn(kz, half_w), bin_fctr);
            if (vbx < 0.0f || vby < 0.0f || vbz < 0.0f || vbx >= (float) kHistPerDim ||
                vby >= (float) kHistPerDim || vbz >= (float) kHistPerDim)
                continue;

            float gx = __fmul_rn(__fmul_rn(0.5f, __fsub_rn(row[x + 1], row[x - 1])), ixu);
            float gy = __fmul_rn(__fmul_rn(0.5f, __fsub_rn(row[x + ys], row[(ptrdiff_t) x - (ptrdiff_t) ys])), iyu);
            float gz = __fmul_rn(__fmul_rn(0.5f, __fsub_rn(row[x + zs], row[(ptrdiff_t) x - (ptrdiff_t) zs])), izu);
            const float w = expf(__fdiv_rn(__fmul_rn(-0.5f, sq), sig2));  // sift.c:1498
            gx = __fmul_rn(gx, w);
            gy = __fmul_rn(gy, w);
            gz = __fmul_rn(gz, w);
            const float rx = __fadd_rn(__fadd_rn(__fmul_rn(r00, gx), __fmul_rn(r10, gy)), __fmul_rn(r20, gz));
            const float ry = __fadd_rn(__fadd_rn(__fmul_rn(r01, gx), __fmul_rn(r11, gy)), __fmul_rn(r21, gz));
            const float rz = __fadd_rn(__fadd_rn(__fmul_rn(r02, gx), __fmul_rn(r12, gy)), __fmul_rn(r22, gz));

            // icos_hist_bin (sift.c:1254-1291)
            const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)), __fmul_rn(rz, rz));
            if ((double) n2 < kBaryEps) continue;
            // Fast selection: the face whose centre direction is closest to the ray. If the
            // ray is comfortably inside that face no other face can pass the reference's
            // (>= -bary_eps) test, so it is also the *first* passing face. Otherwise fall
            // back to the literal first-hit scan so ties on edges/vertices resolve as in
            // the reference.
            int best = 0;
            float bestd = -FLT_MAX;
#pragma unroll
            for (int f = 0; f < kFaces; f++) {
                const float d = fmaf(rx, c_mesh.ctr[f][0], fmaf(ry, c_mesh.ctr[f][1], rz * c_mesh.ctr[f][2]));
                if (d > bestd) { bestd = d; best = f; }
            }
            float bx, by, bz;
            int face = -1;
            if (face_hit(best, rx, ry, rz, bx, by, bz) && bx > 1e-3f && by > 1e-3f && bz > 1e-3f) {
                face = best;
            } else {
                for (int f = 0; f < kFaces; f++) {
                    if (face_hit(f, rx, ry, rz, bx, by, bz)) { face = f; break; }
                }
            }
            if (face < 0) continue;  // "unreachable" in the reference (sift.c:1288-1290)
            const float mag = sqrtf(n2);

            // SIFT3D_desc_acc_interp (sift.c:1318-1397)
            const int bxi = (int) vbx, byi = (int) vby, bzi = (int) vbz;
            const int key = bxi | (byi << 2) | (bzi << 4) | (face << 6);
            if (key != cur_key) {
                S3D_FLUSH();
                cur_key = key;
                cbx = bxi; cby = byi; cbz = bzi;
                ci0 = c_mesh.idx[face][0]; ci1 = c_mesh.idx[face][1]; ci2 = c_mesh.idx[face][2];
#pragma unroll
                for (int c = 0; c < 8; c++) acc[c][0] = acc[c][1] = acc[c][2] = 0.0f;
            }
            const float fx = __fsub_rn(vbx, floorf(vbx));
            const float fy = __fsub_rn(vby, floorf(vby));
            const float fz = __fsub_rn(vbz, floorf(vbz));
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const int ddx = c >> 2, ddy = (c >> 1) & 1, ddz = c & 1;
                const float wgt = __fmul_rn(
                    __fmul_rn(ddx == 0 ? __fsub_rn(1.0f, fx) : fx, ddy == 0 ? __fsub_rn(1.0f, fy) : fy),
                    ddz == 0 ? __fsub_rn(1.0f, fz) : fz);
                const float mw = __fmul_rn(mag, wgt);
                acc[c][0] = __fadd_rn(acc[c][0], __fmul_rn(mw, bx));
                acc[c][1] = __fadd_rn(acc[c][1], __fmul_rn(mw, by));
                acc[c][2] = __fadd_rn(acc[c][2], __fmul_rn(mw, bz));
            }
        }
        S3D_FLUSH();
        cur_key = -1;
    }
#undef S3D_FLUSH
    __syncthreads();
    for (int i = threadIdx.x; i < kDescNumel; i += kDescThreads) {
        const long long q = (long long) (((unsigned long long) hhi[i] << 32) | (unsigned long long) hlo[i]);
        hist[i] = (float) ((double) q * kFixInv);
    }
    __syncthreads();

    // normalise, clamp, normalise (sift.c:1513-1526)
    const float trunc = __fdiv_rn(__fmul_rn(0.2f, 128.0f), (float) kDescNumel);  // sift.c:45
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int pass = 0; pass < 2; pass++) {
        double ssq = 0.0;
        for (int i = threadIdx.x; i < kDescNumel; i += kDescThreads) {
            const float el = hist[i];
            ssq += (double) el * (double) el;
        }
        for (int o = 16; o > 0; o >>= 1) ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
        if (lane == 0) red[warp] = ssq;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int wv = 0; wv < kDescThreads / 32; wv++) t += red[wv];
            norm_sh = sqrt(t) + DBL_EPSILON;
        }
        __syncthreads();
        const float norm_inv = (float) (1.0 / norm_sh);  // sift.c:1423
        for (int i = threadIdx.x; i < kDescNumel; i += kDescThreads) {
            float el = __fmul_rn(hist[i], norm_inv);
            if (pass == 0) el = fminf(el, trunc);
            hist[i] = el;
        }
        __syncthreads();
    }
    float *o = desc_out + (size_t) J.out_index * kDescNumel;
    for (int i = threadIdx.x; i < kDescNumel; i += kDescThreads) o[i] = hist[i];
}

// ---------------------------------------------------------------------------------------
// The window of a keypoint is the sphere of radius 2*sigma intersected with the rotated
// cube of half-width rad/sqrt(2) (only voxels whose rotated coordinates fall into the
// 4x4x4 cell grid contribute, sift.c:1488-1492) -- about a third of the bounding box the
// reference loops over. The exact per-voxel tests of the reference are kept (the closed-form
// x intervals the kernel walks are supersets), so the set of contributing voxels is the
// reference's.
//
// Face lookup: nearest face centre out of the 4 candidates of the direction's octant
// (icosahedral symmetry), confirmed by the reference's own Moller-Trumbore test with a
// margin; anything near an edge falls back to the literal first-hit scan of
// icos_hist_bin so that ties resolve exactly as in the reference.
// ---------------------------------------------------------------------------------------

// Continuous quantities of the cell kernel's per-voxel front end (rotations, the cross and dot
// products of the barycentric solve) may use fused multiply-adds: descriptors are compared
// within a tolerance (cosine >= 0.999), and an FMA only removes an intermediate rounding.
// The quantities the reference DECIDES on exactly once per voxel (the sphere test on sq, the
// literal first-hit face scan of the slow path) keep the reference's unfused operation order.
// -DSIFT3D_DESC_EXACT restores the unfused order everywhere.
#ifdef SIFT3D_DESC_EXACT
constexpr bool kDescFma = false;
#else
constexpr bool kDescFma = true;
#endif
template <bool FMA>
__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    if (FMA) return fmaf(a2, b2, fmaf(a1, b1, __fmul_rn(a0, b0)));
    return __fadd_rn(__fadd_rn(__fmul_rn(a0, b0), __fmul_rn(a1, b1)), __fmul_rn(a2, b2));
}
template <bool FMA>
__device__ __forceinline__ float diff2(float a0, float b0, float a1, float b1) {  // a0*b0 - a1*b1
    if (FMA) return fmaf(a0, b0, -__fmul_rn(a1, b1));
    return __fsub_rn(__fmul_rn(a0, b0), __fmul_rn(a1, b1));
}

// cart2bary on a face held in shared memory: fm = {e1[3], e2[3], t[3], q[3], e2q}
template <bool FMA = false>
__device__ __forceinline__ bool face_hit_s(const float *__restrict__ fm, float gx, float gy,
                                           float gz, float &bx, float &by, float &bz) {
    const float4 m0 = *reinterpret_cast<const float4 *>(fm);       // e1.x e1.y e1.z e2.x
    const float4 m1 = *reinterpret_cast<const float4 *>(fm + 4);   // e2.y e2.z t.x  t.y
    const float4 m2 = *reinterpret_cast<const float4 *>(fm + 8);   // t.z  q.x  q.y  q.z
    const float e2q = fm[12];
    const float e2x = m0.w, e2y = m1.x, e2z = m1.y;
    const float px = diff2<FMA>(gy, e2z, gz, e2y);
    const float py = diff2<FMA>(gz, e2x, gx, e2z);
    const float pz = diff2<FMA>(gx, e2y, gy, e2x);
    const float det = dot3<FMA>(m0.x, px, m0.y, py, m0.z, pz);
    if (fabsf(det) < kBaryEpsF) return false;
    const float det_inv = __frcp_rn(det);  // == 1.0f / det, correctly rounded
    const float tp = dot3<FMA>(m1.z, px, m1.w, py, m2.x, pz);
    const float cq = dot3<FMA>(gx, m2.y, gy, m2.z, gz, m2.w);
    by = __fmul_rn(det_inv, tp);
    bz = __fmul_rn(det_inv, cq);
    bx = __fsub_rn(__fsub_rn(1.0f, by), bz);
    const float k = __fmul_rn(e2q, det_inv);
    if (bx < -kBaryEpsF || by < -kBaryEpsF || bz < -kBaryEpsF || k < 0.0f)
        return false;
    return true;
}

constexpr int kPadDim = kHistPerDim + 1;
constexpr int kPadNumel = kPadDim * kPadDim * kPadDim * kVerts;
constexpr int kFix2LoBits = 12;
constexpr float kFix2Scale = 268435456.0f;  // 2^28 fixed-point units per 1.0
constexpr double kFix2Inv = 1.0 / 268435456.0;

// Carry-free two-word fixed point: q = v * 2^28 as int64, low 12 bits into an unsigned word
// and the (signed) rest into a second word, both with the native 32-bit ATOMS.ADD. The low
// word holds up to 2^20 additions, the high word sums of magnitude < 2^15; integer adds
// commute, so descriptors stay bit-reproducible from run to run.
__device__ __forceinline__ void fix_add2(unsigned *lo, int *hi, int i, float v) {
    const long long q = __float2ll_rn(v * kFix2Scale);
    atomicAdd(lo + i, (unsigned) q & ((1u << kFix2LoBits) - 1u));
    atomicAdd(hi + i, (int) (q >> kFix2LoBits));
}

// ---------------------------------------------------------------------------------------
// v4: cell-stationary accumulation, no atomics on the voxel path.
//
// ncu on v3: the 24 x 2 shared-memory ATOMS.ADD per voxel ran at 2.4 wavefronts per
// instruction (random bank conflicts between lanes that walk far-apart voxels) and the
// run-length flush executed under divergence -- the shared-memory pipe, not arithmetic, set
// the time. v4 turns the loop around the 4x4x4 cell grid:
//   * a warp takes one floor cell (b0, b1, b2) at a time (64 cells, centre-first order,
//     handed out by a shared counter). All 32 lanes then hit the same 8 corner cells, so
//     each lane owns a private copy of the 8 x 12 bins the cell can touch, laid out
//     [bin][lane]: bank = lane, every LDS/STS is one conflict-free wavefront and the
//     accumulation is plain fp32 LDS + FMA + STS in a fixed order;
//   * the voxels of a cell come from closed-form x intervals: 32 rows of the cell's
//     bounding box at a time (one per lane: chord of the sphere /\ three slabs of the rotated
//     cell, with a margin), a warp scan, and lane t of step k takes flattened entry 32k + t --
//     neighbouring lanes are neighbouring x, so the six gradient loads coalesce;
//   * the reference's exact per-voxel tests are kept and a voxel is accepted only by the cell
//     its own (int) vb indices name, so every voxel is counted exactly once whatever the
//     margins;
//   * when the cell is done the 96 x 32 partial sums are reduced (rotated reads, conflict-free)
//     and added to the block histogram in 2^-28 fixed point (96 x 2 integer atomics per cell
//     visit; integer adds commute and the per-cell sums have a fixed order, so descriptors
//     are reproducible run to run).
// No window-size limit: nothing is tabulated per row.
// ---------------------------------------------------------------------------------------
constexpr int kCellBins = 8 * kVerts;  // bins one floor cell can touch
__constant__ unsigned char c_cell_order[64];

static int upload_cell_order() {
    static PerDeviceOnce once;
    return once.run([&]() -> int {
    unsigned char order[64];
    int key[64];
    for (int c = 0; c < 64; c++) {
        const int b0 = c & 3, b1 = (c >> 2) & 3, b2 = c >> 4;
        const int d0 = 2 * b0 - 3, d1 = 2 * b1 - 3, d2 = 2 * b2 - 3;  // twice the centre offset
        key[c] = (d0 * d0 + d1 * d1 + d2 * d2) * 64 + c;
        order[c] = (unsigned char) c;
    }
    for (int i = 0; i < 64; i++)
        for (int j = i + 1; j < 64; j++)
            if (key[order[j]] < key[order[i]]) {
                const unsigned char t = order[i];
                order[i] = order[j];
                order[j] = t;
            }
    S3D_CK(cudaMemcpyToSymbol(c_cell_order, order, sizeof(order)));
    return 0;
    });
}

// `split` CTAs (a thread-block cluster) share one keypoint when there are fewer keypoints than CTA
// slots (the CLI's 100 strongest, small octaves of a slab): CTA r of the cluster takes the cells
// r, r + split, ... of the centre-first order, rank 0 adds the others' block histograms through
// distributed shared memory. The histograms are integers and every cell's sum is formed by one
// warp in a fixed order, so the descriptor has the same bits for every split.
__global__ void __launch_bounds__(kDescThreads)
descriptor_cell_kernel(const DescJob *__restrict__ jobs, int n, float *__restrict__ desc_out, int split) {
    extern __shared__ __align__(16) float wacc_all[];  // [warps][kCellBins][32]
    __shared__ unsigned hlo[kPadNumel];
    __shared__ int hhi[kPadNumel];
    __shared__ __align__(16) float smesh[kFaces * 16];
    // fast path, by (octant, which of its 4 candidate faces): rows of [v0 v1 v2]^-1 with the
    // histogram bin of the row's vertex in .w -- one lookup instead of octant -> face -> rows / bins
    __shared__ __align__(16) float sfast[32 * 12];
    __shared__ int sidx[kFaces * 3];
    __shared__ unsigned char slut[32];
    __shared__ double red[kDescThreads / 32];
    __shared__ double norm_sh;
    __shared__ int next_cell;
    __shared__ unsigned char rowlane_all[kDescThreads];  // per warp: lane of the k-th non-empty row
    const int ki = blockIdx.x / split, part = blockIdx.x - ki * split;
    if (ki >= n) return;  // (uniform over a cluster)
    const DescJob J = jobs[ki];
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const unsigned full = 0xffffffffu;
    float *__restrict__ wacc = wacc_all + warp * (kCellBins * 32);
    unsigned char *rowlane = rowlane_all + warp * 32;
    for (int i = tid; i < kPadNumel; i += kDescThreads) {
        hlo[i] = 0u;
        hhi[i] = 0;
    }
    if (tid < kFaces) {
        float *fm = smesh + tid * 16;
        for (int k = 0; k < 3; k++) {
            fm[k] = c_mesh.e1[tid][k];
            fm[3 + k] = c_mesh.e2[tid][k];
            fm[6 + k] = c_mesh.t[tid][k];
            fm[9 + k] = c_mesh.q[tid][k];
            sidx[tid * 3 + k] = c_mesh.idx[tid][k];
        }
        fm[12] = c_mesh.e2q[tid];
        fm[13] = fm[14] = fm[15] = 0.0f;
    }
    if (tid < 32) {
        const int f = c_mesh.lut[tid >> 2][tid & 3];
        slut[tid] = (unsigned char) f;
        for (int k = 0; k < 12; k++)
            sfast[tid * 12 + k] = (k & 3) == 3 ? __int_as_float(c_mesh.idx[f][k >> 2] * 32) : c_mesh.minv[f][k];
    }
    if (tid == 0) next_cell = part;

    const float *__restrict__ img = J.img;
    const int nx = J.nx, ny = J.ny, nz = J.nz;
    const float uxf = J.ux, uyf = J.uy, uzf = J.uz;
    const float cx = J.cx, cy = J.cy, cz = J.cz;
    // sift.c:1453-1458
    const float sigma = J.sigma;
    const float rad = (float) (kDescRadFctr * (double) sigma);
    const float half_w = (float) ((double) rad / sqrt(2.0));
    const float desc_w = __fmul_rn(2.0f, half_w);
    const float hist_w = __fdiv_rn(desc_w, (float) kHistPerDim);
    const float bin_fctr = __fdiv_rn(1.0f, hist_w);
    const float rad2 = __fmul_rn(rad, rad);
    const float sig2 = __fmul_rn(sigma, sigma);
    const float nhs = -0.5f / sig2;
    const float nhs2 = nhs * 1.4426950408889634f;  // exp(x) = 2^(x log2 e)
    const float ixu = __fdiv_rn(1.0f, uxf), iyu = __fdiv_rn(1.0f, uyf), izu = __fdiv_rn(1.0f, uzf);
    const float hixu = 0.5f * ixu, hiyu = 0.5f * iyu, hizu = 0.5f * izu;
    const float hw_bf = half_w * bin_fctr;  // == 2: (k + half_w) * bin_fctr as one FMA

    // Window (sift.c:91-99), all-float here because win_radius is a float.
    const int x0 = (int) fmaxf(floorf(__fsub_rn(cx, __fdiv_rn(rad, uxf))), 1.0f);
    const int x1 = (int) fminf(ceilf(__fadd_rn(cx, __fdiv_rn(rad, uxf))), (float) (nx - 2));
    const int y0 = (int) fmaxf(floorf(__fsub_rn(cy, __fdiv_rn(rad, uyf))), 1.0f);
    const int y1 = (int) fminf(ceilf(__fadd_rn(cy, __fdiv_rn(rad, uyf))), (float) (ny - 2));
    const int z0 = (int) fmaxf(floorf(__fsub_rn(cz, __fdiv_rn(rad, uzf))), 1.0f);
    const int z1 = (int) fminf(ceilf(__fadd_rn(cz, __fdiv_rn(rad, uzf))), (float) (nz - 2));
    const size_t ys = (size_t) nx, zs = (size_t) nx * ny;

    // Rt = R^T (sift.c:1462-1464): Rt[i][j] = R[j][i]
    const float r00 = J.R[0], r01 = J.R[1], r02 = J.R[2];
    const float r10 = J.R[3], r11 = J.R[4], r12 = J.R[5];
    const float r20 = J.R[6], r21 = J.R[7], r22 = J.R[8];
    const float avx = r00 * uxf, avy = r01 * uxf, avz = r02 * uxf;  // d(vkp)/dx per voxel
    const float cmar = hist_w * 1.0e-3f + 1.0e-6f;                  // slab margin
    const int cyi = (int) floorf(cy), czi = (int) floorf(cz);
    const float cyf = cy - (float) cyi, czf = cz - (float) czi;
    const bool av_ok[3] = {fabsf(avx) > 1e-6f, fabsf(avy) > 1e-6f, fabsf(avz) > 1e-6f};
    const float av_inv[3] = {av_ok[0] ? 1.0f / avx : 0.0f, av_ok[1] ? 1.0f / avy : 0.0f,
                             av_ok[2] ? 1.0f / avz : 0.0f};
    __syncthreads();

    for (;;) {
        int ci = 0;
        if (lane == 0) ci = atomicAdd(&next_cell, split);
        ci = __shfl_sync(full, ci, 0);
        if (ci >= 64) break;
        const int cell = c_cell_order[ci];
        const int b0 = cell & 3, b1 = (cell >> 2) & 3, b2 = cell >> 4;
        const float fb0 = (float) b0, fb1 = (float) b1, fb2 = (float) b2;
        // rotated-frame slab of the cell along each axis, and its centre
        const float lo0 = (float) b0 * hist_w - half_w, lo1 = (float) b1 * hist_w - half_w,
                    lo2 = (float) b2 * hist_w - half_w;
        const float kc0 = lo0 + 0.5f * hist_w, kc1 = lo1 + 0.5f * hist_w, kc2 = lo2 + 0.5f * hist_w;
        // bounding rows: vim = R * vkp; half extents of the rotated cube along y and z
        const float dyc = r10 * kc0 + r11 * kc1 + r12 * kc2;
        const float dzc = r20 * kc0 + r21 * kc1 + r22 * kc2;
        const float ey = 0.5f * hist_w * (fabsf(r10) + fabsf(r11) + fabsf(r12));
        const float ez = 0.5f * hist_w * (fabsf(r20) + fabsf(r21) + fabsf(r22));
        // (integer part of the centre + offset: the row box, and with it the order in which
        // voxels are summed, does not change when the volume is translated by whole voxels --
        // a z-slab of a taller volume gives the same bits as the volume itself)
        const int ya = max(y0, cyi + (int) floorf(cyf + (dyc - ey) * iyu) - 2);
        const int yb = min(y1, cyi + (int) ceilf(cyf + (dyc + ey) * iyu) + 2);
        const int za = max(z0, czi + (int) floorf(czf + (dzc - ez) * izu) - 2);
        const int zb = min(z1, czi + (int) ceilf(czf + (dzc + ez) * izu) + 2);
        const int wyc = yb - ya + 1, wzc = zb - za + 1;
        if (wyc <= 0 || wzc <= 0 || x1 < x0) continue;
        const int nrows = wyc * wzc;
        const float inv_wyc = 1.0f / (float) wyc;
#pragma unroll 8
        for (int i = 0; i < kCellBins / 4; i++)
            *reinterpret_cast<float4 *>(wacc + i * 128 + lane * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        bool any = false;

        // Rows are taken 32 at a time (one per lane): x interval of row /\ sphere /\ cell, then
        // an exclusive scan -> a set of up to 32 segments holding `T` flattened entries. Steps
        // consume 32 entries; when the current set runs short the step is topped up from the
        // next set, so only the last step of a cell can be partial.
        auto load_set = [&](int batch, int &pre, int &xa, int &ry, int &T) {
            const int rr = batch * 32 + lane;
            int len = 0, rz = za;
            xa = x0; ry = ya;
            if (rr < nrows) {
                // rr / wyc by a float reciprocal and one correction step (exact for any operands
                // below 2^23; an integer division is ~20 instructions)
                int q = (int) (((float) rr + 0.5f) * inv_wyc);
                int rem_y = rr - q * wyc;
                if (rem_y < 0) { q--; rem_y += wyc; }
                else if (rem_y >= wyc) { q++; rem_y -= wyc; }
                if (nrows >= (1 << 22)) { q = rr / wyc; rem_y = rr - q * wyc; }  // (warp-uniform; never in practice)
                rz = za + q;
                ry = ya + rem_y;
                const float dy = ((float) ry - cy) * uyf, dz = ((float) rz - cz) * uzf;
                const float rem = rad2 * 1.0001f - dy * dy - dz * dz;
                if (rem >= 0.0f) {
                    const float half = sqrtf(rem) * ixu + 0.01f;
                    float lo = cx - half, hi = cx + half;
                    bool empty = false;
                    const float bv[3] = {r10 * dy + r20 * dz, r11 * dy + r21 * dz, r12 * dy + r22 * dz};
                    const float sl[3] = {lo0, lo1, lo2};
#pragma unroll
                    for (int i = 0; i < 3; i++) {
                        const float a = sl[i] - cmar - bv[i], b = sl[i] + hist_w + cmar - bv[i];
                        if (av_ok[i]) {
                            const float t1 = a * av_inv[i], t2 = b * av_inv[i];
                            lo = fmaxf(lo, cx + fminf(t1, t2) - 0.01f);
                            hi = fminf(hi, cx + fmaxf(t1, t2) + 0.01f);
                        } else if (a > 0.0f || b < 0.0f) {
                            empty = true;
                        }
                    }
                    if (!empty) {
                        xa = max(x0, (int) ceilf(lo));
                        const int xb = min(x1, (int) floorf(hi));
                        len = max(0, xb - xa + 1);
                    }
                }
            }
            int incl = len;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(full, incl, o);
                if (lane >= o) incl += v;
            }
            T = __shfl_sync(full, incl, 31);
            // compact the non-empty rows to lanes 0 .. K-1 (in order): their starts are then
            // strictly increasing and the row of an entry is a population count (next_entries)
            const unsigned nb = __ballot_sync(full, len > 0);
            if (len > 0) rowlane[__popc(nb & ((1u << lane) - 1u))] = (unsigned char) lane;
            __syncwarp();
            const int src = rowlane[lane];
            __syncwarp();
            const int ex = incl - len;
            const int p = __shfl_sync(full, ex, src);
            const int xm = __shfl_sync(full, xa - ex, src);
            const int pk = __shfl_sync(full, ((rz - za) << 16) | (ry - ya), src);
            const bool held = lane < __popc(nb);
            pre = held ? p : 0x7fffffff;
            xa = xm;   // x of entry t of this row: xa + t
            ry = pk;   // (z - za) << 16 | (y - ya)
        };
        const int nbatch = (nrows + 31) >> 5;
        int bi = 0;
        int c_pre, c_xm, c_yz, c_T, c_off = 0;
        int n_pre = 0x7fffffff, n_xm = 0, n_yz = 0, n_T = 0;
        bool have_next = false;
        load_set(bi++, c_pre, c_xm, c_yz, c_T);
        const unsigned le_mask = 0xffffffffu >> (31 - lane);  // bits 0 .. lane
        // the next 32 flattened entries: one voxel per lane (warp-uniform control flow)
        auto next_entries = [&](int &x, int &y, int &z, bool &active) -> bool {
            for (;;) {
                if (c_T - c_off > 0) break;
                if (have_next) {
                    c_pre = n_pre; c_xm = n_xm; c_yz = n_yz; c_T = n_T; c_off = 0;
                    have_next = false;
                } else if (bi < nbatch) {
                    load_set(bi++, c_pre, c_xm, c_yz, c_T);
                    c_off = 0;
                } else {
                    x = x0; y = ya; z = za; active = false;
                    return false;
                }
            }
            const int avail = c_T - c_off;
            const int fill_a = min(avail, 32);
            bool mixed = false;
            if (avail < 32 && (have_next || bi < nbatch)) {
                if (!have_next) {
                    load_set(bi++, n_pre, n_xm, n_yz, n_T);
                    have_next = true;
                }
                mixed = true;
            }
            const int take_b = mixed ? min(32 - fill_a, n_T) : 0;
            const bool from_a = lane < fill_a;
            const bool from_b = !from_a && lane - fill_a < take_b;
            // Row of entry c_off + lane: the rows that start at or before c_off (a ballot) plus the
            // starts inside the step up to this lane (one bit per start, OR-reduced: REDUX) -- the
            // held rows are non-empty, so the starts are distinct.
            const int sa = c_pre - c_off;
            const unsigned bits_a = __reduce_or_sync(full, (unsigned) (sa - 1) < 31u ? 1u << sa : 0u);
            const int ja = __popc(__ballot_sync(full, sa <= 0)) - 1 + __popc(bits_a & le_mask);
            int xm = __shfl_sync(full, c_xm, ja);
            int yz = __shfl_sync(full, c_yz, ja);
            int t = c_off + lane;
            if (take_b > 0) {  // warp-uniform: the step is topped up from the next set
                const int sb = n_pre < 32 ? n_pre + fill_a : 32;
                const unsigned bits_b = __reduce_or_sync(full, sb < 32 ? 1u << sb : 0u);
                const int jb = __popc(bits_b & le_mask) - 1;
                const int xmb = __shfl_sync(full, n_xm, jb);
                const int yzb = __shfl_sync(full, n_yz, jb);
                if (from_b) { xm = xmb; yz = yzb; t = lane - fill_a; }
            }
            x = xm + t;
            y = ya + (yz & 0xffff);
            z = za + (yz >> 16);
            c_off += fill_a;
            if (mixed && c_off >= c_T) {  // the current set is used up: the next one takes over
                c_pre = n_pre; c_xm = n_xm; c_yz = n_yz; c_T = n_T; c_off = take_b;
                have_next = false;
            }
            active = from_a || from_b;
            if (!active) { x = x0; y = ya; z = za; }  // a legal address for the unused lanes
            return true;
        };

        // One voxel up to its face and magnitudes, WITHOUT branches: two voxels per lane are
        // evaluated back to back so that their (long, serial) dependency chains interleave --
        // with 4 warps per scheduler a single chain leaves ~45 % of the issue slots empty.
        struct Vox {
            bool valid, slow;
            float rx, ry, rz, n2, mag;
            float b0, b1, b2;
            int i0, i1, i2;  // float offsets of the three vertex bins in a lane's accumulators
            float fx, fy, fz;
        };
        auto eval_fast = [&](int x, int y, int z, bool active) -> Vox {
            Vox v;
            const float dy = __fmul_rn(__fsub_rn((float) y, cy), uyf);
            const float dz = __fmul_rn(__fsub_rn((float) z, cz), uzf);
            const float dx = __fmul_rn(__fsub_rn((float) x, cx), uxf);
            const float sq = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            bool ok = active && !(sq > rad2);
            // vkp = Rt * vim (SIFT3D_MUL_MAT_RM_CVEC, immacros.h:326-342)
            const float kx = dot3<kDescFma>(r00, dx, r10, dy, r20, dz);
            const float ky = dot3<kDescFma>(r01, dx, r11, dy, r21, dz);
            const float kz = dot3<kDescFma>(r02, dx, r12, dy, r22, dz);
            const float vbx = kDescFma ? fmaf(kx, bin_fctr, hw_bf) : __fmul_rn(__fadd_rn(kx, half_w), bin_fctr);
            const float vby = kDescFma ? fmaf(ky, bin_fctr, hw_bf) : __fmul_rn(__fadd_rn(ky, half_w), bin_fctr);
            const float vbz = kDescFma ? fmaf(kz, bin_fctr, hw_bf) : __fmul_rn(__fadd_rn(kz, half_w), bin_fctr);
            // sift.c:1481-1486 rejects vb outside [0, 4) and bins by (int) vb; the voxel belongs to the
            // cell its own indices name (the interval margins overlap). For vb >= 0 the cast is the
            // floor, for vb < 0 the floor is <= -1 and matches no cell: one compare per axis.
            const float flx = floorf(vbx), fly = floorf(vby), flz = floorf(vbz);
            ok = ok && flx == fb0 && fly == fb1 && flz == fb2;
            v.fx = __fsub_rn(vbx, flx);
            v.fy = __fsub_rn(vby, fly);
            v.fz = __fsub_rn(vbz, flz);

            const float *__restrict__ row = img + (size_t) z * zs + (size_t) y * ys;
            // 0.5 * (a - b) / u: the halving is exact, so (a - b) * (0.5 / u) has the same bits
            float gx = __fmul_rn(__fsub_rn(row[x + 1], row[x - 1]), hixu);
            float gy = __fmul_rn(__fsub_rn(row[x + ys], row[(ptrdiff_t) x - (ptrdiff_t) ys]), hiyu);
            float gz = __fmul_rn(__fsub_rn(row[x + zs], row[(ptrdiff_t) x - (ptrdiff_t) zs]), hizu);
            float w;  // exp(-0.5 * sq / sigma^2), sift.c:1498: one multiply and MUFU.EX2
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(w) : "f"(sq * nhs2));
            gx = __fmul_rn(gx, w);
            gy = __fmul_rn(gy, w);
            gz = __fmul_rn(gz, w);
            const float rx = dot3<kDescFma>(r00, gx, r10, gy, r20, gz);
            const float ry2 = dot3<kDescFma>(r01, gx, r11, gy, r21, gz);
            const float rz2 = dot3<kDescFma>(r02, gx, r12, gy, r22, gz);
            // icos_hist_bin (sift.c:1254-1291)
            const float n2 = dot3<kDescFma>(rx, rx, ry2, ry2, rz2, rz2);
            ok = ok && !(n2 < kBaryEpsF);
            // nearest face centre among the 4 candidates of this octant (selection only)
            const float ax = fabsf(rx), ay = fabsf(ry2), az = fabsf(rz2);
            const float kA = 2.6180339887f, kB = 1.6180339887f, kC = 4.2360679775f;  // 1+phi, phi, 2phi+1
            const float d0 = kA * (ax + ay + az);
            const float d1 = fmaf(kB, ax, kC * az);
            const float d2 = fmaf(kC, ay, kB * az);
            const float d3 = fmaf(kC, ax, kB * ay);
            // first maximum of (d0, d1, d2, d3), as a two-level tree (same ties as a serial scan)
            const int w01 = d1 > d0 ? 1 : 0, w23 = d3 > d2 ? 3 : 2;
            const int which = fmaxf(d2, d3) > fmaxf(d0, d1) ? w23 : w01;
            const int oct = (rx < 0.0f ? 1 : 0) | (ry2 < 0.0f ? 2 : 0) | (rz2 < 0.0f ? 4 : 0);
            float bx, by, bz;
            bool hit;
            if (kDescFma) {
                // b' = [v0 v1 v2]^-1 r, b = b' / sum(b'): 9 multiply-adds and one reciprocal instead
                // of the cross products of Moeller-Trumbore (same point, a few ulp apart); a
                // direction within 1e-5 of an edge goes to the literal scan as before
                const float *fe = sfast + (oct * 4 + which) * 12;
                const float4 m0 = *reinterpret_cast<const float4 *>(fe);
                const float4 m1 = *reinterpret_cast<const float4 *>(fe + 4);
                const float4 m2 = *reinterpret_cast<const float4 *>(fe + 8);
                v.i0 = __float_as_int(m0.w); v.i1 = __float_as_int(m1.w); v.i2 = __float_as_int(m2.w);
                const float p0 = dot3<true>(m0.x, rx, m0.y, ry2, m0.z, rz2);
                const float p1 = dot3<true>(m1.x, rx, m1.y, ry2, m1.z, rz2);
                const float p2 = dot3<true>(m2.x, rx, m2.y, ry2, m2.z, rz2);
                const float sum = p0 + p1 + p2;
                const float inv = __fdividef(1.0f, sum);  // MUFU.RCP: the barycentric weights are tolerance quantities
                bx = p0 * inv; by = p1 * inv; bz = p2 * inv;
                hit = sum > 0.0f && bx > 1e-5f && by > 1e-5f && bz > 1e-5f;
            } else {
                const int best = slut[oct * 4 + which];
                hit = face_hit_s<false>(smesh + best * 16, rx, ry2, rz2, bx, by, bz) && bx > 1e-5f &&
                      by > 1e-5f && bz > 1e-5f;
                v.i0 = sidx[best * 3] * 32; v.i1 = sidx[best * 3 + 1] * 32; v.i2 = sidx[best * 3 + 2] * 32;
            }
            v.valid = ok;
            v.slow = ok && !hit;  // near an edge: the literal first-hit scan decides
            v.rx = rx; v.ry = ry2; v.rz = rz2; v.n2 = n2;
            // (here, not in accumulate(): off the LDS -> FMA -> STS chain) MUFU.SQRT, 1 ulp
            asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(v.mag) : "f"(n2));
            v.b0 = bx; v.b1 = by; v.b2 = bz;
            return v;
        };
        auto eval_slow = [&](Vox &v) {
            float bx, by, bz;
            int face = -1;
            for (int f = 0; f < kFaces; f++) {
                if (face_hit_s(smesh + f * 16, v.rx, v.ry, v.rz, bx, by, bz)) { face = f; break; }
            }
            if (face < 0) {  // "unreachable" in the reference (sift.c:1288-1290)
                v.valid = false;
            } else {
                v.i0 = sidx[face * 3] * 32; v.i1 = sidx[face * 3 + 1] * 32; v.i2 = sidx[face * 3 + 2] * 32;
                v.b0 = bx; v.b1 = by; v.b2 = bz;
            }
        };
        // SIFT3D_desc_acc_interp (sift.c:1318-1397) into this lane's copy of the cell's
        // 8 x 12 bins: corner c = (ddx, ddy, ddz), bin (c, vertex)
        auto accumulate = [&](const Vox &v) {
            const float mag = v.mag;
            const float m3[3] = {__fmul_rn(mag, v.b0), __fmul_rn(mag, v.b1), __fmul_rn(mag, v.b2)};
            // a lane's copy of the cell's bins: [vertex][ddx][lane][ddy ddz] -- the four (ddy, ddz)
            // corners of a (vertex, ddx) are one 16-byte word, so a voxel is 6 LDS.128 + 6 STS.128
            // (each four conflict-free wavefronts) instead of 24 + 24 32-bit accesses
            float4 *a4[3];
#pragma unroll
            for (int k = 0; k < 3; k++)
                a4[k] = reinterpret_cast<float4 *>(wacc + (k == 0 ? v.i0 : (k == 1 ? v.i1 : v.i2)) * 8 + lane * 4);
            const float wx0 = __fsub_rn(1.0f, v.fx), wy0 = __fsub_rn(1.0f, v.fy), wz0 = __fsub_rn(1.0f, v.fz);
            const float4 yz = make_float4(__fmul_rn(wy0, wz0), __fmul_rn(wy0, v.fz), __fmul_rn(v.fy, wz0),
                                          __fmul_rn(v.fy, v.fz));
            // all loads first, then the FMAs, then the stores: written bin by bin the compiler
            // must assume the three vertex columns alias and serialises LDS -> FFMA -> STS
            float4 acc[3][2];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                acc[k][0] = a4[k][0];
                acc[k][1] = a4[k][32];
            }
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const float m0 = __fmul_rn(m3[k], wx0), m1 = __fmul_rn(m3[k], v.fx);
                acc[k][0].x = fmaf(m0, yz.x, acc[k][0].x); acc[k][0].y = fmaf(m0, yz.y, acc[k][0].y);
                acc[k][0].z = fmaf(m0, yz.z, acc[k][0].z); acc[k][0].w = fmaf(m0, yz.w, acc[k][0].w);
                acc[k][1].x = fmaf(m1, yz.x, acc[k][1].x); acc[k][1].y = fmaf(m1, yz.y, acc[k][1].y);
                acc[k][1].z = fmaf(m1, yz.z, acc[k][1].z); acc[k][1].w = fmaf(m1, yz.w, acc[k][1].w);
            }
#pragma unroll
            for (int k = 0; k < 3; k++) {
                a4[k][0] = acc[k][0];
                a4[k][32] = acc[k][1];
            }
        };

        for (;;) {
            int xA, yA, zA, xB, yB, zB;
            bool actA, actB;
            if (!next_entries(xA, yA, zA, actA)) break;
            next_entries(xB, yB, zB, actB);
            any = true;
            Vox va = eval_fast(xA, yA, zA, actA);
            Vox vb = eval_fast(xB, yB, zB, actB);
            if (va.slow) eval_slow(va);
            if (vb.slow) eval_slow(vb);
            if (va.valid) accumulate(va);
            if (vb.valid) accumulate(vb);
        }
        __syncwarp();
        if (any) {
            // ---- reduce the 32 lane copies of each bin and add to the block histogram
            const int hb = (b0 + b1 * kPadDim + b2 * kPadDim * kPadDim) * kVerts;
#pragma unroll
            for (int q = 0; q < kCellBins / 32; q++) {
                const int t = q * 32 + lane;
                const int g = t >> 2, j = t & 3;  // g = vertex * 2 + ddx, j = ddy * 2 + ddz
                float sum = 0.0f;
                // lanes of one j start 8 banks apart ((lane >> 2) * 4 + j): conflict-free
#pragma unroll 8
                for (int i = 0; i < 32; i++)
                    sum = __fadd_rn(sum, wacc[g * 128 + (((lane >> 2) + i) & 31) * 4 + j]);
                const int v = g >> 1, c = ((g & 1) << 2) | j;
                const int off = ((c >> 2) + ((c >> 1) & 1) * kPadDim + (c & 1) * kPadDim * kPadDim) * kVerts;
                if (sum != 0.0f) fix_add2(hlo, hhi, hb + off + v, sum);
            }
        }
        __syncwarp();
    }
    __syncthreads();
    if (split > 1) {
        namespace cg = cooperative_groups;
        cg::cluster_group cluster = cg::this_cluster();
        cluster.sync();  // every CTA of the keypoint has finished its cells
        if (part == 0) {
            for (int r = 1; r < split; r++) {
                const unsigned *rlo = cluster.map_shared_rank(hlo, r);
                const int *rhi = cluster.map_shared_rank(hhi, r);
                for (int i = tid; i < kPadNumel; i += kDescThreads) {
                    hlo[i] += rlo[i];
                    hhi[i] += rhi[i];
                }
            }
        }
        cluster.sync();  // the others' shared memory stays alive until rank 0 has read it
        if (part != 0) return;
        __syncthreads();
    }
    float *hist = wacc_all;  // the lane copies are dead: reuse (keeps two CTAs per SM in 228 KB)
    for (int i = tid; i < kDescNumel; i += kDescThreads) {
        const int bin = i % kVerts, cell = i / kVerts;
        const int hx = cell & 3, hy = (cell >> 2) & 3, hz = cell >> 4;
        const int j = (hx + hy * kPadDim + hz * kPadDim * kPadDim) * kVerts + bin;
        const long long q = (long long) hhi[j] * (1ll << kFix2LoBits) + (long long) hlo[j];
        hist[i] = (float) ((double) q * kFix2Inv);
    }
    __syncthreads();

    // normalise, clamp, normalise (sift.c:1513-1526)
    const float trunc = __fdiv_rn(__fmul_rn(0.2f, 128.0f), (float) kDescNumel);  // sift.c:45
    for (int pass = 0; pass < 2; pass++) {
        double ssq = 0.0;
        for (int i = tid; i < kDescNumel; i += kDescThreads) {
            const float el = hist[i];
            ssq += (double) el * (double) el;
        }
        for (int o = 16; o > 0; o >>= 1) ssq += __shfl_xor_sync(full, ssq, o);
        if (lane == 0) red[warp] = ssq;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int wv = 0; wv < kDescThreads / 32; wv++) t += red[wv];
            norm_sh = sqrt(t) + DBL_EPSILON;
        }
        __syncthreads();
        const float norm_inv = (float) (1.0 / norm_sh);  // sift.c:1423
        for (int i = tid; i < kDescNumel; i += kDescThreads) {
            float el = __fmul_rn(hist[i], norm_inv);
            if (pass == 0) el = fminf(el, trunc);
            hist[i] = el;
        }
        __syncthreads();
    }
    float *o = desc_out + (size_t) J.out_index * kDescNumel;
    for (int i = tid; i < kDescNumel; i += kDescThreads) o[i] = hist[i];
}

__global__ void __launch_bounds__(256)
pull_words_kernel(const unsigned *__restrict__ src, unsigned *__restrict__ dst, size_t n) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

int launch_pull_words(cudaStream_t st, const void *pinned_src, void *dst, size_t n, uint64_t *launches) {
    if (n == 0) return 0;
    static_assert(sizeof(DescJob) % 4 == 0, "DescJob is copied in 32-bit words");
    pull_words_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>((const unsigned *) pinned_src,
                                                                   (unsigned *) dst, n);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

// Host: upper bound of the (y, z) rows of a job's window (the kernel clamps the window to
// the volume, which only makes it smaller).
int desc_job_rows(const DescJob &J) {
    const double rad = 2.0 * (double) J.sigma * 1.001;
    const double wy = 2.0 * ceil(rad / (double) J.uy) + 3.0, wz = 2.0 * ceil(rad / (double) J.uz) + 3.0;
    const double wyc = wy < (double) J.ny ? wy : (double) J.ny;
    const double wzc = wz < (double) J.nz ? wz : (double) J.nz;
    const double rows = wyc * wzc;
    return rows > 2.0e9 ? 2000000000 : (int) rows;
}

int launch_descriptors(cudaStream_t st, const DescJob *d_jobs, int n, int max_rows, float *d_desc,
                       uint64_t *launches) {
    if (n <= 0) return 0;
    // $SIFT3D_CUDA_DESC_ROWS: the literal reference-loop kernel (diagnostics)
    static const bool force_rows = getenv("SIFT3D_CUDA_DESC_ROWS") != nullptr;
    if (!force_rows) {
        const size_t dyn4 = (size_t) (kDescThreads / 32) * kCellBins * 32 * sizeof(float);
        static PerDeviceOnce attr4;
        if (attr4.run([&]() -> int {
                if (upload_cell_order()) return -1;
                S3D_CK(cudaFuncSetAttribute(descriptor_cell_kernel,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int) dyn4));
                return 0;
            }))
            return -1;
        // Few keypoints: the kernel lasts as long as its largest keypoint (0.3 - 0.45 ms) however
        // empty the device is, so keypoints are split over 4 CTAs up to one wave of keypoints (2 CTA
        // slots per SM) and over 2 up to two waves ($SIFT3D_CUDA_DESC_SPLIT=1|2|4 forces a value)
        const char *se = getenv("SIFT3D_CUDA_DESC_SPLIT");  // (read per call: tests switch it)
        const int split_env = se ? atoi(se) : 0;
        int split = split_env == 1 || split_env == 2 || split_env == 4 ? split_env
                    : (n <= 2 * kNumSMs ? 4 : (n <= 4 * kNumSMs ? 2 : 1));
        if (split == 1) {
            descriptor_cell_kernel<<<(unsigned) n, kDescThreads, dyn4, st>>>(d_jobs, n, d_desc, 1);
        } else {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned) n * (unsigned) split, 1, 1);
            cfg.blockDim = dim3(kDescThreads, 1, 1);
            cfg.dynamicSmemBytes = dyn4;
            cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = (unsigned) split;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            S3D_CK(cudaLaunchKernelEx(&cfg, descriptor_cell_kernel, d_jobs, n, d_desc, split));
        }
        if (launches) *launches += 1;
        S3D_CK(cudaGetLastError());
        return 0;
    }
    (void) max_rows;
    descriptor_rows_kernel<<<(unsigned) n, kDescThreads, 0, st>>>(d_jobs, n, d_desc);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

}  // namespace s3d
