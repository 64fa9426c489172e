/* nifti1.c -- self-contained NIfTI-1 / Analyze 7.5 reader and NIfTI-1 writer.
 *
 * The reference reads volumes only through nifticlib (nifti.c:52-167, a thin wrapper over
 * nifti_image_read), which is an optional dependency and absent in most builds -- then
 * sift3d_read_image fails for every path (nifti.c:16-31). This file replaces the dependency:
 * it restates what the wrapper takes from a nifti_image, so that sift3d_read_image works
 * everywhere with the result the reference + nifticlib would give:
 *
 *   dims    the array's dimensionality is the last dim[i] > 1 (nifti.c:66-71); more than 4
 *           is an error (:75-79); 4-D is a 3-D volume with nt channels (:96)
 *   units   ux,uy,uz = dx,dy,dz = pixdim[1..3] (nifti.c:89-91); nifticlib replaces zero or
 *           non-finite pixdim by 1 when it loads the header
 *   values  (float) ((double) raw * slope + inter), slope 0 -> 1 (nifti.c:100-111), for
 *           the ten integer / real datatypes of nifti.c:114-150; anything else is an error
 *   layout  file index x + nx*(y + ny*(z + nz*c)) (nifti.c:43-46) -> channel-fastest
 *           c + nc*(x + nx*(y + ny*z)) (imutil.c:520-533)
 *
 * Files: single-file .nii / .nii.gz (magic "n+1"), and header/data pairs .hdr + .img
 * (optionally .gz; magic "ni1", or none for plain Analyze 7.5, where slope and intercept
 * do not exist). Either byte order. zlib's gz* functions read plain files transparently.
 * Our own code: nifticlib is not part of the reference tree and nothing is copied from it;
 * the format facts are those of the public NIfTI-1 specification (nifti1.h field table).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include "nifti1.h"

#define NII_HDR_SIZE 348

/* byte offsets of the NIfTI-1 header fields used here */
enum {
    OFF_SIZEOF_HDR = 0,  /* int32, must be 348 */
    OFF_DIM = 40,        /* int16[8] */
    OFF_DATATYPE = 70,   /* int16 */
    OFF_BITPIX = 72,     /* int16 */
    OFF_PIXDIM = 76,     /* float32[8] */
    OFF_VOX_OFFSET = 108,/* float32 */
    OFF_SCL_SLOPE = 112, /* float32 */
    OFF_SCL_INTER = 116, /* float32 */
    OFF_MAGIC = 344      /* char[4] */
};

#define NII_ERR(...) fprintf(stderr, __VA_ARGS__)

static uint16_t rd16(const unsigned char *p, int swap) {
    return swap ? (uint16_t) (p[0] << 8 | p[1]) : (uint16_t) (p[1] << 8 | p[0]);
}

static uint32_t rd32(const unsigned char *p, int swap) {
    return swap ? ((uint32_t) p[0] << 24 | (uint32_t) p[1] << 16 | (uint32_t) p[2] << 8 | p[3])
                : ((uint32_t) p[3] << 24 | (uint32_t) p[2] << 16 | (uint32_t) p[1] << 8 | p[0]);
}

static uint64_t rd64(const unsigned char *p, int swap) {
    const uint64_t a = rd32(p, swap), b = rd32(p + 4, swap);
    return swap ? (a << 32 | b) : (b << 32 | a);
}

static float rdf32(const unsigned char *p, int swap) {
    const uint32_t u = rd32(p, swap);
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static double rdf64(const unsigned char *p, int swap) {
    const uint64_t u = rd64(p, swap);
    double d;
    memcpy(&d, &u, 8);
    return d;
}

static int ends_with(const char *s, const char *suffix) {
    const size_t n = strlen(s), m = strlen(suffix);
    return n >= m && strcmp(s + n - m, suffix) == 0;
}

/* Read exactly n bytes (gzread takes an unsigned count). */
static int gz_read_all(gzFile f, void *dst, size_t n) {
    unsigned char *p = (unsigned char *) dst;
    while (n > 0) {
        const unsigned chunk = n > (1u << 30) ? (1u << 30) : (unsigned) n;
        const int got = gzread(f, p, chunk);
        if (got <= 0)
            return -1;
        p += got;
        n -= (size_t) got;
    }
    return 0;
}

static int gz_skip(gzFile f, size_t n) {
    unsigned char buf[4096];
    while (n > 0) {
        const unsigned chunk = n > sizeof(buf) ? (unsigned) sizeof(buf) : (unsigned) n;
        if (gzread(f, buf, chunk) != (int) chunk)
            return -1;
        n -= chunk;
    }
    return 0;
}

/* Size in bytes of one element of a supported datatype, 0 if unsupported. */
static int dtype_size(int datatype) {
    switch (datatype) {
    case 2: case 256:             return 1; /* uint8, int8 */
    case 4: case 512:             return 2; /* int16, uint16 */
    case 8: case 768: case 16:    return 4; /* int32, uint32, float32 */
    case 1024: case 1280: case 64: return 8; /* int64, uint64, float64 */
    default:                      return 0;
    }
}

static double element(const unsigned char *p, int datatype, int swap) {
    switch (datatype) {
    case 2:    return (double) p[0];
    case 256:  return (double) (int8_t) p[0];
    case 4:    return (double) (int16_t) rd16(p, swap);
    case 512:  return (double) rd16(p, swap);
    case 8:    return (double) (int32_t) rd32(p, swap);
    case 768:  return (double) rd32(p, swap);
    case 1024: return (double) (int64_t) rd64(p, swap);
    case 1280: return (double) rd64(p, swap);
    case 16:   return (double) rdf32(p, swap);
    default:   return rdf64(p, swap); /* 64 */
    }
}

/* Header file of `path`: itself for .nii(.gz) / .hdr(.gz); the matching .hdr for .img(.gz).
 * Returns 1 if the data live in a separate file (written to img_path). */
static int pair_paths(const char *path, char *hdr_path, char *img_path, size_t cap) {
    const int gz = ends_with(path, ".gz");
    const size_t n = strlen(path) - (gz ? 3 : 0);
    if (n + 8 > cap)
        return -1;
    if (n >= 4 && (!strncmp(path + n - 4, ".img", 4) || !strncmp(path + n - 4, ".hdr", 4))) {
        memcpy(hdr_path, path, n - 4);
        memcpy(img_path, path, n - 4);
        sprintf(hdr_path + n - 4, ".hdr%s", gz ? ".gz" : "");
        sprintf(img_path + n - 4, ".img%s", gz ? ".gz" : "");
        return 1;
    }
    strcpy(hdr_path, path);
    strcpy(img_path, path);
    return 0;
}

int nifti1_read(const char *path, nifti1_info *info, nifti1_alloc_fn alloc, void *user,
                float **data) {
    unsigned char h[NII_HDR_SIZE];
    char hdr_path[4096], img_path[4096];
    gzFile f = NULL;
    unsigned char *raw = NULL;
    float *out = NULL;
    int pair, swap, ndim, i, dimensionality, datatype, esz, is_nifti;
    int dim[8];
    double pix[8], slope, inter;
    size_t nvox, offset, x, y, z, c, nx, ny, nz, nc;

    *data = NULL;
    if ((pair = pair_paths(path, hdr_path, img_path, sizeof(hdr_path))) < 0) {
        NII_ERR("read_nii: path too long: %s\n", path);
        return -1;
    }
    if ((f = gzopen(hdr_path, "rb")) == NULL || gz_read_all(f, h, NII_HDR_SIZE)) {
        NII_ERR("read_nii: failure loading file %s\n", hdr_path);
        goto fail;
    }
    /* byte order: sizeof_hdr is 348 in the file's own order */
    if (rd32(h + OFF_SIZEOF_HDR, 0) == NII_HDR_SIZE)
        swap = 0;
    else if (rd32(h + OFF_SIZEOF_HDR, 1) == NII_HDR_SIZE)
        swap = 1;
    else {
        NII_ERR("read_nii: %s is not a NIfTI-1 or Analyze 7.5 header\n", hdr_path);
        goto fail;
    }
    /* "n+1" = data in the same file, "ni1" = separate .img; no magic = Analyze 7.5 */
    is_nifti = h[OFF_MAGIC] == 'n' && (h[OFF_MAGIC + 1] == '+' || h[OFF_MAGIC + 1] == 'i') &&
               h[OFF_MAGIC + 2] >= '1' && h[OFF_MAGIC + 2] <= '9' && h[OFF_MAGIC + 3] == 0;
    if (is_nifti && h[OFF_MAGIC + 1] == '+' && pair) {
        /* a single-file image that happens to be called .img / .hdr */
        strcpy(img_path, hdr_path);
        pair = 0;
    } else if (is_nifti && h[OFF_MAGIC + 1] == 'i' && !pair) {
        NII_ERR("read_nii: %s holds a header only; pass the .img or .hdr path\n", path);
        goto fail;
    } else if (!is_nifti && !pair) {
        NII_ERR("read_nii: %s has no NIfTI magic\n", path);
        goto fail;
    }

    ndim = (int16_t) rd16(h + OFF_DIM, swap);
    if (ndim < 1 || ndim > 7) {
        NII_ERR("read_nii: file %s has bad dim[0] = %d\n", path, ndim);
        goto fail;
    }
    for (i = 1; i <= 7; i++) {
        dim[i] = (int16_t) rd16(h + OFF_DIM + 2 * i, swap);
        pix[i] = (double) rdf32(h + OFF_PIXDIM + 4 * i, swap);
        if (i > ndim || dim[i] <= 0)
            dim[i] = 1; /* unused or bad sizes count as 1 */
        if (i <= ndim && (pix[i] == 0.0 || !isfinite(pix[i])))
            pix[i] = 1.0; /* as nifticlib repairs the header on load */
    }
    /* dimensionality = last dimension greater than 1 (nifti.c:66-71) */
    for (dimensionality = ndim; dimensionality > 0; dimensionality--)
        if (dim[dimensionality] > 1)
            break;
    if (dimensionality > 4) {
        NII_ERR("read_nii: file %s has unsupported dimensionality %d\n", path, dimensionality);
        goto fail;
    }
    datatype = (int16_t) rd16(h + OFF_DATATYPE, swap);
    if ((esz = dtype_size(datatype)) == 0) {
        NII_ERR("read_nii: unsupported datatype %d \n", datatype);
        goto fail;
    }
    slope = is_nifti ? (double) rdf32(h + OFF_SCL_SLOPE, swap) : 0.0;
    inter = is_nifti ? (double) rdf32(h + OFF_SCL_INTER, swap) : 0.0;
    if (!isfinite(slope)) slope = 0.0; /* nifticlib zeroes non-finite header floats */
    if (!isfinite(inter)) inter = 0.0;
    if (slope == 0.0) slope = 1.0;     /* ill-formatted: ignore (nifti.c:100-102) */

    nx = (size_t) dim[1];
    ny = (size_t) dim[2];
    nz = (size_t) dim[3];
    nc = dimensionality == 4 ? (size_t) dim[4] : 1;
    nvox = nx * ny * nz * nc;
    {
        const float vo = rdf32(h + OFF_VOX_OFFSET, swap);
        offset = isfinite(vo) && vo > 0.0f ? (size_t) vo : 0;
        if (!pair && offset < NII_HDR_SIZE + 4)
            offset = NII_HDR_SIZE + 4; /* header + extension flag */
    }
    if (pair) {
        gzclose(f);
        if ((f = gzopen(img_path, "rb")) == NULL) {
            NII_ERR("read_nii: failure loading file %s\n", img_path);
            goto fail;
        }
        if (gz_skip(f, offset))
            goto short_file;
    } else if (gz_skip(f, offset - NII_HDR_SIZE)) {
        goto short_file;
    }
    if ((raw = (unsigned char *) malloc(nvox * (size_t) esz)) == NULL ||
        (out = (float *) alloc(nvox * sizeof(float), user)) == NULL) {
        NII_ERR("read_nii: out of memory for %zu voxels\n", nvox);
        goto fail;
    }
    if (gz_read_all(f, raw, nvox * (size_t) esz))
        goto short_file;
    gzclose(f);
    f = NULL;

    for (c = 0; c < nc; c++)
        for (z = 0; z < nz; z++)
            for (y = 0; y < ny; y++) {
                const unsigned char *src = raw + (size_t) esz * (nx * (y + ny * (z + nz * c)));
                float *dst = out + c + nc * (nx * (y + ny * z));
                for (x = 0; x < nx; x++)
                    dst[nc * x] = (float) (element(src + (size_t) esz * x, datatype, swap) * slope + inter);
            }
    free(raw);
    info->nx = (int) nx;
    info->ny = (int) ny;
    info->nz = (int) nz;
    info->nc = (int) nc;
    info->ux = pix[1];
    info->uy = pix[2];
    info->uz = pix[3];
    *data = out;
    return 0;

short_file:
    NII_ERR("read_nii: file %s ends before the voxel data do\n", pair ? img_path : hdr_path);
fail:
    if (f != NULL)
        gzclose(f);
    free(raw);
    /* `out`, if any, belongs to the caller's allocator: it is reported so it can be freed */
    *data = out;
    return -1;
}

static void wr16(unsigned char *p, int v) {
    p[0] = (unsigned char) (v & 0xff);
    p[1] = (unsigned char) ((v >> 8) & 0xff);
}

static void wr32(unsigned char *p, uint32_t v) {
    p[0] = (unsigned char) (v & 0xff);
    p[1] = (unsigned char) ((v >> 8) & 0xff);
    p[2] = (unsigned char) ((v >> 16) & 0xff);
    p[3] = (unsigned char) ((v >> 24) & 0xff);
}

static void wrf32(unsigned char *p, float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    wr32(p, u);
}

/* write_nii (nifti.c:170-222): float32, 3-D or 4-D (channels as nt), no scaling. */
int nifti1_write(const char *path, const nifti1_info *info, const float *data) {
    unsigned char h[NII_HDR_SIZE + 4];
    const int gz = ends_with(path, ".gz");
    const size_t nx = (size_t) info->nx, ny = (size_t) info->ny, nz = (size_t) info->nz,
                 nc = (size_t) info->nc;
    const int multi = info->nc > 1;
    gzFile f;
    float *plane;
    size_t x, y, z, c;
    int ok = 1;

    if (!(ends_with(path, ".nii") || ends_with(path, ".nii.gz"))) {
        NII_ERR("write_nii: %s: only .nii and .nii.gz are written\n", path);
        return -1;
    }
    if (info->nx > 32767 || info->ny > 32767 || info->nz > 32767 || info->nc > 32767) {
        NII_ERR("write_nii: dimensions exceed the NIfTI-1 limit of 32767\n");
        return -1;
    }
    memset(h, 0, sizeof(h));
    wr32(h + OFF_SIZEOF_HDR, NII_HDR_SIZE);
    wr16(h + OFF_DIM, multi ? 4 : 3);
    wr16(h + OFF_DIM + 2, info->nx);
    wr16(h + OFF_DIM + 4, info->ny);
    wr16(h + OFF_DIM + 6, info->nz);
    wr16(h + OFF_DIM + 8, multi ? info->nc : 1);
    wr16(h + OFF_DIM + 10, 1);
    wr16(h + OFF_DIM + 12, 1);
    wr16(h + OFF_DIM + 14, 1);
    wr16(h + OFF_DATATYPE, 16); /* NIFTI_TYPE_FLOAT32 */
    wr16(h + OFF_BITPIX, 32);
    wrf32(h + OFF_PIXDIM, 1.0f); /* qfac */
    wrf32(h + OFF_PIXDIM + 4, (float) info->ux);
    wrf32(h + OFF_PIXDIM + 8, (float) info->uy);
    wrf32(h + OFF_PIXDIM + 12, (float) info->uz);
    wrf32(h + OFF_PIXDIM + 16, multi ? 0.0f : 1.0f); /* channels have no size (nifti.c:193) */
    wrf32(h + OFF_VOX_OFFSET, (float) (NII_HDR_SIZE + 4));
    wrf32(h + OFF_SCL_SLOPE, 1.0f);
    wrf32(h + OFF_SCL_INTER, 0.0f);
    memcpy(h + OFF_MAGIC, "n+1", 4);

    /* level 1 keeps .gz writing fast; plain files pass through uncompressed ("T") */
    if ((f = gzopen(path, gz ? "wb1" : "wbT")) == NULL) {
        NII_ERR("write_nii: cannot open %s\n", path);
        return -1;
    }
    if ((plane = (float *) malloc(nx * ny * sizeof(float))) == NULL) {
        gzclose(f);
        return -1;
    }
    ok = gzwrite(f, h, sizeof(h)) == (int) sizeof(h);
    for (c = 0; ok && c < nc; c++)
        for (z = 0; ok && z < nz; z++) {
            for (y = 0; y < ny; y++)
                for (x = 0; x < nx; x++) {
                    unsigned char b[4];
                    wrf32(b, data[c + nc * (x + nx * (y + ny * z))]);
                    memcpy(plane + x + nx * y, b, 4);
                }
            {
                const unsigned char *p = (const unsigned char *) plane;
                size_t left = nx * ny * sizeof(float);
                while (ok && left > 0) {
                    const unsigned chunk = left > (1u << 30) ? (1u << 30) : (unsigned) left;
                    ok = gzwrite(f, p, chunk) == (int) chunk;
                    p += chunk;
                    left -= chunk;
                }
            }
        }
    free(plane);
    if (gzclose(f) != Z_OK)
        ok = 0;
    if (!ok)
        NII_ERR("write_nii: error writing %s\n", path);
    return ok ? 0 : -1;
}
