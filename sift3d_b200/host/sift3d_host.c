/* sift3d_host.c -- host side of the drop-in library (libsift3D.so): the reference's
 * 27 public functions (include/sift3d.h) in plain C on top of the CUDA C ABI
 * (include/sift3d_cuda.h). No compute happens here: this file owns the opaque
 * objects, validates parameters like the reference does, derives the pyramid
 * geometry / scales / Gaussian taps with the reference's own double->float
 * expressions, and converts device results into the caller-visible stores.
 *
 * Reference behaviour mirrored here (paths under /root/reference/sift3d):
 *   parameter setters + ranges          sift.c:499-565
 *   octave count, pyramid geometry      sift.c:427-475, imutil.c:1464-1554
 *   level scales                        imutil.c:1564-1599
 *   Gaussian filter design              imutil.c:1267-1343, 1360-1409
 *   "re-plan only if dims changed"      sift.c:651-656
 *   survivor compaction (strength bug)  sift.c:1148-1166, copy_Keypoint :372-384
 *   verify_keys                         sift.c:1171-1212
 *   stores -> matrices / CSV            sift.c:1644-1830, imutil.c:405-479
 *   sort_by_strength                    sift.c:1832-1900
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include "../../include/sift3d.h"
#include "../../include/sift3d_cuda.h"
#include "nifti1.h"

#define ERR(...) fprintf(stderr, __VA_ARGS__)

#define DESC_NUMEL SIFT3D_CUDA_DESC_NUMEL
#define IM_NDIMS 3

/* Defaults (sift.c:31-35) */
static const double peak_thresh_default = 0.1;
static const int num_kp_levels_default = 3;
static const double corner_thresh_default = 0.4;
static const double sigma_n_default = 1.15;
static const double sigma0_default = 1.6;
#define GAUSS_WIDTH_FCTR 3.0 /* imutil.c:1264-1266 */

/* ---- object layouts (private; the reference's public headers keep them opaque) ---- */

struct _sift3d_image {
    float *data; /* pinned host memory */
    size_t size; /* voxels * channels */
    int nx, ny, nz, nc;
    double ux, uy, uz;
    void *dev;   /* optional HBM-resident copy (sift3d_image_make_resident) */
    int dev_id;
    sift3d_cuda_upload *up; /* asynchronous upload ticket (sift3d_image_prefetch) */
    int up_pending;         /* the ticket holds a transfer no detect has consumed yet */
};

struct _sift3d_mat_rm {
    void *data;
    size_t size; /* bytes */
    int num_cols, num_rows;
    sift3d_mat_type type;
};

typedef struct {
    double xd, yd, zd, sd;
    int o, s;
    float strength;
    float R[9];
} keypoint_rec;

struct _sift3d_keypoint_store {
    keypoint_rec *buf;
    size_t num, cap;
    int nx, ny, nz; /* dims of the first octave (sift.c:756-759) */
};

struct _sift3d_descriptor_store {
    float *desc;  /* num * 768, page-locked */
    double *xyzs; /* num * 4: base-octave x, y, z and scale */
    size_t num, cap;
    int nx, ny, nz;
};

struct _sift3d_detector {
    sift3d_cuda_ctx *ctx;
    double peak_thresh, corner_thresh, sigma_n, sigma0;
    int num_kp_levels;
    /* geometry of the image the pyramid was last sized for */
    int have_dims;
    int nx, ny, nz;
    int plan_valid;    /* device plan matches dims + parameters */
    int have_gpyr;     /* a detect has completed (detector_has_gpyr, sift.c:1544-1549) */
    sift3d_cuda_keypoint *scratch;
    size_t scratch_cap;
    /* one volume over several devices (sift3d_detect_describe_sharded): cached per geometry */
    sift3d_cuda_multi *multi;
    double multi_key[32];
};

/* ---- matrices (imutil.c:134-290, 1676-1710) ------------------------------------------ */

static size_t mat_type_size(sift3d_mat_type t) {
    switch (t) {
    case SIFT3D_DOUBLE: return sizeof(double);
    case SIFT3D_FLOAT: return sizeof(float);
    case SIFT3D_INT: return sizeof(int);
    }
    return 0;
}

/* resize_Mat_rm (imutil.c:226-279) */
static int mat_resize(sift3d_mat_rm *m, int rows, int cols, sift3d_mat_type type) {
    const size_t ts = mat_type_size(type);
    size_t total;
    if (ts == 0) {
        ERR("resize_Mat_rm: unknown type! \n");
        return SIFT3D_FAILURE;
    }
    m->num_rows = rows;
    m->num_cols = cols;
    m->type = type;
    total = ts * (size_t) rows * (size_t) cols;
    if (total == m->size)
        return SIFT3D_SUCCESS;
    if (total == 0) {
        free(m->data);
        m->data = NULL;
        m->size = 0;
        return SIFT3D_SUCCESS;
    }
    {
        void *p = realloc(m->data, total);
        if (p == NULL) {
            free(m->data);
            m->data = NULL;
            m->size = 0;
            return SIFT3D_FAILURE;
        }
        m->data = p;
        m->size = total;
    }
    return SIFT3D_SUCCESS;
}

sift3d_mat_rm *sift3d_make_mat_rm(void) {
    sift3d_mat_rm *m = (sift3d_mat_rm *) calloc(1, sizeof(*m));
    if (m == NULL)
        return NULL;
    m->type = SIFT3D_FLOAT; /* imutil.c:1678 */
    return m;
}

void sift3d_free_mat_rm(sift3d_mat_rm *m) {
    if (m == NULL)
        return;
    free(m->data);
    free(m);
}

void *sift3d_mat_rm_data(sift3d_mat_rm *m) { return m->data; }

void sift3d_mat_rm_dimensions(const sift3d_mat_rm *m, int *num_cols, int *num_rows) {
    if (num_cols != NULL)
        *num_cols = m->num_cols;
    if (num_rows != NULL)
        *num_rows = m->num_rows;
}

sift3d_mat_type sift3d_mat_rm_type(const sift3d_mat_rm *m) { return m->type; }

/* write_Mat_rm (imutil.c:405-479): "%f" for double and float, "%d" for int, gzip when
 * the extension is "gz". */
static const char *file_ext(const char *path) {
    const char *name = strrchr(path, '/');
    const char *dot;
    name = name == NULL ? path : name;
    dot = strrchr(name, '.');
    return dot == NULL || dot == name ? "" : dot + 1;
}

static int mat_write(const char *path, const sift3d_mat_rm *m) {
    const int compress = strcmp(file_ext(path), "gz") == 0;
    FILE *file = NULL;
    gzFile gz = NULL;
    int i, j;

    if (compress) {
        if ((gz = gzopen(path, "w")) == Z_NULL)
            return SIFT3D_FAILURE;
    } else if ((file = fopen(path, "w")) == NULL) {
        return SIFT3D_FAILURE;
    }
    for (i = 0; i < m->num_rows; i++) {
        for (j = 0; j < m->num_cols; j++) {
            const char delim = j < m->num_cols - 1 ? ',' : '\n';
            const size_t k = (size_t) j + (size_t) i * (size_t) m->num_cols;
            char buf[512];
            int len;
            switch (m->type) {
            case SIFT3D_DOUBLE: len = snprintf(buf, sizeof(buf), "%f", ((double *) m->data)[k]); break;
            case SIFT3D_FLOAT: len = snprintf(buf, sizeof(buf), "%f", ((float *) m->data)[k]); break;
            default: len = snprintf(buf, sizeof(buf), "%d", ((int *) m->data)[k]); break;
            }
            if (len < 0 || len >= (int) sizeof(buf) - 1)
                goto fail;
            buf[len++] = delim;
            if (compress) {
                if (gzwrite(gz, buf, (unsigned) len) != len)
                    goto fail;
            } else if (fwrite(buf, 1, (size_t) len, file) != (size_t) len) {
                goto fail;
            }
        }
    }
    if (compress)
        return gzclose(gz) == Z_OK ? SIFT3D_SUCCESS : SIFT3D_FAILURE;
    if (ferror(file))
        goto fail;
    fclose(file);
    return SIFT3D_SUCCESS;
fail:
    if (compress)
        gzclose(gz);
    else
        fclose(file);
    return SIFT3D_FAILURE;
}

/* ---- images (imutil.c:493-586, 1639-1674) -------------------------------------------- */

sift3d_image *sift3d_make_image(const int nx, const int ny, const int nz, const int nc) {
    sift3d_image *im;
    const int dims[3] = {nx, ny, nz};
    int i;

    /* im_resize checks (imutil.c:560-575) */
    for (i = 0; i < IM_NDIMS; i++) {
        if (dims[i] > 0)
            continue;
        ERR("im_resize: invalid dimension %d: %d \n", i, dims[i]);
        return NULL;
    }
    if (nc < 1) {
        ERR("im_resize: invalid number of channels: %d \n", nc);
        return NULL;
    }
    if ((im = (sift3d_image *) calloc(1, sizeof(*im))) == NULL)
        return NULL;
    im->nx = nx;
    im->ny = ny;
    im->nz = nz;
    im->nc = nc;
    im->ux = im->uy = im->uz = 1.0; /* init_im, imutil.c:1249-1251 */
    im->size = (size_t) nx * (size_t) ny * (size_t) nz * (size_t) nc;
    im->data = (float *) sift3d_cuda_host_alloc(im->size * sizeof(float));
    if (im->data == NULL) {
        ERR("sift3d_make_image: cannot allocate page-locked memory for %zu voxels "
            "(is a CUDA device present?)\n", im->size);
        free(im);
        return NULL;
    }
    memset(im->data, 0, im->size * sizeof(float)); /* im_zero, imutil.c:507 */
    return im;
}

void sift3d_free_image(sift3d_image *im) {
    if (im == NULL)
        return;
    if (im->dev != NULL)
        sift3d_cuda_device_free(im->dev_id, im->dev);
    sift3d_cuda_upload_free(im->up);
    sift3d_cuda_host_free(im->data);
    free(im);
}

/* Extension: snapshot the current contents into device memory; later detects on a
 * detector bound to the same device read that copy instead of uploading the host
 * buffer. Call again after changing the host data. device < 0 drops the copy. */
int sift3d_image_make_resident(sift3d_image *im, int device) {
    if (im == NULL)
        return SIFT3D_FAILURE;
    if (im->dev != NULL) {
        sift3d_cuda_device_free(im->dev_id, im->dev);
        im->dev = NULL;
    }
    if (device < 0)
        return SIFT3D_SUCCESS;
    im->dev = sift3d_cuda_device_upload(device, im->data, im->size);
    im->dev_id = device;
    return im->dev != NULL ? SIFT3D_SUCCESS : SIFT3D_FAILURE;
}

/* Extension: start copying the current contents to device `device` and return at once. The
 * next sift3d_detect_keypoints on a detector of that device consumes the copy (its stream
 * waits for the transfer; the host does not), so the upload of volume i+1 hides behind the
 * kernels of volume i. The host buffer must stay unchanged until that detect. */
int sift3d_image_prefetch(sift3d_image *im, int device) {
    if (im == NULL || device < 0)
        return SIFT3D_FAILURE;
    if (im->dev != NULL && im->dev_id == device)
        return SIFT3D_SUCCESS; /* a resident copy is what the detect will read: nothing to upload */
    im->up = sift3d_cuda_upload_begin(device, im->data, im->size, im->up);
    im->up_pending = im->up != NULL;
    return im->up != NULL ? SIFT3D_SUCCESS : SIFT3D_FAILURE;
}

/* Extension: block until the transfer started by sift3d_image_prefetch has completed. */
int sift3d_image_prefetch_wait(sift3d_image *im) {
    if (im == NULL || im->up == NULL)
        return SIFT3D_FAILURE;
    return sift3d_cuda_upload_sync(im->up) ? SIFT3D_FAILURE : SIFT3D_SUCCESS;
}

static void *pinned_alloc(size_t bytes, void *user) {
    (void) user;
    return sift3d_cuda_host_alloc(bytes);
}

static void *plain_alloc(size_t bytes, void *user) {
    (void) user;
    return malloc(bytes ? bytes : 1);
}

/* im_read (imutil.c:352-372) + read_nii (nifti.c:52-167). The reference needs nifticlib for
 * this and fails on every path without it (nifti.c:16-31); here the format is read by
 * nifti1.c, so .nii, .nii.gz and .hdr/.img pairs always work. */
sift3d_image *sift3d_read_image(const char *path) {
    const char *ext;
    sift3d_image *im;
    nifti1_info info;
    float *data = NULL;

    if (path == NULL)
        return NULL;
    ext = file_ext(path);
    /* im_get_format (imutil.c:318-331) */
    if (strcmp(ext, "img") && strcmp(ext, "gz") && strcmp(ext, "nii") && strcmp(ext, "hdr")) {
        ERR("im_read: unrecognized file extension from file %s \n", path);
        return NULL;
    }
    if (nifti1_read(path, &info, pinned_alloc, NULL, &data)) {
        sift3d_cuda_host_free(data);
        return NULL;
    }
    if ((im = (sift3d_image *) calloc(1, sizeof(*im))) == NULL) {
        sift3d_cuda_host_free(data);
        return NULL;
    }
    im->nx = info.nx;
    im->ny = info.ny;
    im->nz = info.nz;
    im->nc = info.nc;
    /* nifti.c:89-91 stores pixdim as is; a negative spacing would make every tap offset
     * negative, so its magnitude is used */
    im->ux = fabs(info.ux);
    im->uy = fabs(info.uy);
    im->uz = fabs(info.uz);
    im->size = (size_t) info.nx * (size_t) info.ny * (size_t) info.nz * (size_t) info.nc;
    im->data = data;
    return im;
}

/* Extension: im_write (imutil.c:385-402) + write_nii (nifti.c:170-222): float32 NIfTI-1,
 * .nii or .nii.gz, units in pixdim. */
int sift3d_write_image(const char *path, const sift3d_image *im) {
    nifti1_info info;
    if (path == NULL || im == NULL)
        return SIFT3D_FAILURE;
    info.nx = im->nx;
    info.ny = im->ny;
    info.nz = im->nz;
    info.nc = im->nc;
    info.ux = im->ux;
    info.uy = im->uy;
    info.uz = im->uz;
    return nifti1_write(path, &info, im->data) ? SIFT3D_FAILURE : SIFT3D_SUCCESS;
}

/* Extension (tools and tests on machines without a GPU): the reader without the page-locked
 * image object. dims = nx, ny, nz, nc; *data is malloc'ed, release with sift3d_nifti_free. */
int sift3d_nifti_read_raw(const char *path, int dims[4], double units[3], float **data) {
    nifti1_info info;
    if (path == NULL || dims == NULL || units == NULL || data == NULL)
        return SIFT3D_FAILURE;
    if (nifti1_read(path, &info, plain_alloc, NULL, data)) {
        free(*data);
        *data = NULL;
        return SIFT3D_FAILURE;
    }
    dims[0] = info.nx; dims[1] = info.ny; dims[2] = info.nz; dims[3] = info.nc;
    units[0] = info.ux; units[1] = info.uy; units[2] = info.uz;
    return SIFT3D_SUCCESS;
}

int sift3d_nifti_write_raw(const char *path, const int dims[4], const double units[3],
                           const float *data) {
    nifti1_info info;
    if (path == NULL || dims == NULL || units == NULL || data == NULL)
        return SIFT3D_FAILURE;
    info.nx = dims[0]; info.ny = dims[1]; info.nz = dims[2]; info.nc = dims[3];
    info.ux = units[0]; info.uy = units[1]; info.uz = units[2];
    return nifti1_write(path, &info, data) ? SIFT3D_FAILURE : SIFT3D_SUCCESS;
}

void sift3d_nifti_free(float *data) { free(data); }

float *sift3d_image_data(const sift3d_image *im) { return im->data; }

int sift3d_image_set_units(sift3d_image *im, double ux, double uy, double uz) {
    if (im == NULL || !(ux > 0) || !(uy > 0) || !(uz > 0)) {
        ERR("sift3d_image_set_units: units must be positive\n");
        return SIFT3D_FAILURE;
    }
    im->ux = ux;
    im->uy = uy;
    im->uz = uz;
    return SIFT3D_SUCCESS;
}

void sift3d_image_get_units(const sift3d_image *im, double *ux, double *uy, double *uz) {
    if (ux) *ux = im->ux;
    if (uy) *uy = im->uy;
    if (uz) *uz = im->uz;
}

void sift3d_image_dimensions(const sift3d_image *im, int *nx, int *ny, int *nz, int *nc) {
    if (nx) *nx = im->nx;
    if (ny) *ny = im->ny;
    if (nz) *nz = im->nz;
    if (nc) *nc = im->nc;
}

/* ---- pyramid geometry, scales and filters --------------------------------------------- */

/* init_Gauss_filter (imutil.c:1267-1319) */
static int design_gauss(const double sigma, sift3d_cuda_filter *f) {
    const int half_width = sigma > 0 ? (((int) ceil(sigma * GAUSS_WIDTH_FCTR)) > 1
                                            ? (int) ceil(sigma * GAUSS_WIDTH_FCTR) : 1)
                                     : 1;
    const int width = 2 * half_width + 1;
    float acc;
    int i;

    if (width > SIFT3D_CUDA_MAX_TAPS) {
        ERR("init_Gauss_filter: filter width %d exceeds the supported maximum %d \n", width,
            SIFT3D_CUDA_MAX_TAPS);
        return SIFT3D_FAILURE;
    }
    acc = 0;
    for (i = 0; i < width; i++) {
        double x = (double) i - half_width;
        x /= sigma + DBL_EPSILON;
        f->taps[i] = (float) exp(-0.5 * x * x);
        acc += f->taps[i];
    }
    for (i = 0; i < width; i++)
        f->taps[i] /= acc;
    f->width = width;
    return SIFT3D_SUCCESS;
}

/* init_Gauss_incremental_filter (imutil.c:1322-1343) */
static int design_gauss_incremental(const double s_cur, const double s_next,
                                    sift3d_cuda_filter *f) {
    if (s_cur > s_next) {
        ERR("init_Gauss_incremental_filter: s_cur (%f) > s_next (%f) \n", s_cur, s_next);
        return SIFT3D_FAILURE;
    }
    return design_gauss(sqrt(s_next * s_next - s_cur * s_cur), f);
}

/* Scale of level (o, s): set_scales_Pyramid (imutil.c:1578-1579) */
static double level_scale(const sift3d_detector *d, int o, int s) {
    return d->sigma0 * pow(2.0, o + (double) s / d->num_kp_levels);
}

/* resize_SIFT3D (sift.c:427-475) + make_gss (imutil.c:1360-1409) written into a plan
 * description. `base_octave` > 0 describes the part of a pyramid that starts at that octave
 * (nx, ny, nz and units are then those of that octave); `num_octaves` 0 derives the count
 * from the dimensions as the reference does. */
static int fill_plan(const sift3d_detector *d, int nx, int ny, int nz, const double units[3],
                     int base_octave, int num_octaves, sift3d_cuda_plan_desc *P) {
    const int first_level = -1;
    int min_dim, last_octave, num_levels, o, s, ret;

    min_dim = nx < ny ? nx : ny;
    min_dim = min_dim < nz ? min_dim : nz;
    /* The minimum size of a pyramid level is 8 in any dimension (sift.c:441-452) */
    last_octave = (int) log2((double) min_dim) - 3;
    if (last_octave < 0) {
        ERR("resize_SIFT3D: input image is too small: must have at least 8 voxels in each "
            "dimension \n");
        return SIFT3D_FAILURE;
    }
    if (num_octaves <= 0)
        num_octaves = last_octave + 1;
    else if (num_octaves > last_octave + 1) {
        ERR("resize_SIFT3D: %d octaves do not fit a %d x %d x %d image \n", num_octaves, nx, ny,
            nz);
        return SIFT3D_FAILURE;
    }
    num_levels = d->num_kp_levels + 3; /* sift.c:434-435 */
    if (d->num_kp_levels < 1 || d->num_kp_levels + 2 > SIFT3D_CUDA_MAX_FILTERS ||
        num_octaves > SIFT3D_CUDA_MAX_OCTAVES || base_octave < 0) {
        ERR("resize_SIFT3D: unsupported pyramid shape (%d keypoint levels, %d octaves) \n",
            d->num_kp_levels, num_octaves);
        return SIFT3D_FAILURE;
    }
    /* set_scales_Pyramid refuses a nominal scale above the first level (imutil.c:1582-1588) */
    if (level_scale(d, 0, first_level) < d->sigma_n) {
        ERR("set_scales_Pyramid: sigma_n too large for these settings. Max allowed: %f \n",
            level_scale(d, 0, first_level) - DBL_EPSILON);
        return SIFT3D_FAILURE;
    }
    memset(P, 0, sizeof(*P));
    P->nx = nx;
    P->ny = ny;
    P->nz = nz;
    P->units[0] = units[0];
    P->units[1] = units[1];
    P->units[2] = units[2];
    P->num_octaves = num_octaves;
    P->num_kp_levels = d->num_kp_levels;
    for (o = 0; o < num_octaves; o++)
        for (s = first_level; s < first_level + num_levels; s++)
            P->scales[o][s - first_level] = level_scale(d, o + base_octave, s);
    ret = design_gauss_incremental(d->sigma_n, level_scale(d, 0, first_level), &P->first);
    for (s = first_level; !ret && s < first_level + num_levels - 1; s++)
        ret = design_gauss_incremental(level_scale(d, 0, s), level_scale(d, 0, s + 1),
                                       &P->octave[s - first_level]);
    return ret;
}

/* Plan the detector's own context for its current dims. */
static int detector_plan(sift3d_detector *d, const double units[3]) {
    sift3d_cuda_plan_desc *P;
    int ret;

    if ((P = (sift3d_cuda_plan_desc *) calloc(1, sizeof(*P))) == NULL)
        return SIFT3D_FAILURE;
    ret = fill_plan(d, d->nx, d->ny, d->nz, units, 0, 0, P);
    if (!ret)
        ret = sift3d_cuda_plan(d->ctx, P) ? SIFT3D_FAILURE : SIFT3D_SUCCESS;
    free(P);
    return ret;
}

/* Extension (multi-GPU slab harness): the plan this detector's parameters give for an
 * arbitrary part of a pyramid, for a caller that drives sift3d_cuda.h contexts itself. */
int sift3d_detector_fill_plan(const sift3d_detector *d, int nx, int ny, int nz, double ux,
                              double uy, double uz, int base_octave, int num_octaves,
                              void *plan_desc) {
    const double units[3] = {ux, uy, uz};
    if (d == NULL || plan_desc == NULL)
        return SIFT3D_FAILURE;
    return fill_plan(d, nx, ny, nz, units, base_octave, num_octaves,
                     (sift3d_cuda_plan_desc *) plan_desc);
}

/* ---- detector ---------------------------------------------------------------------- */

sift3d_detector *sift3d_make_detector_on_device(int device) {
    sift3d_detector *d = (sift3d_detector *) calloc(1, sizeof(*d));
    if (d == NULL)
        return NULL;
    d->peak_thresh = peak_thresh_default;
    d->corner_thresh = corner_thresh_default;
    d->num_kp_levels = num_kp_levels_default;
    d->sigma_n = sigma_n_default;
    d->sigma0 = sigma0_default;
    d->ctx = sift3d_cuda_ctx_create(device);
    if (d->ctx == NULL) {
        ERR("sift3d_make_detector: no usable CUDA device; this build has no CPU path\n");
        free(d);
        return NULL;
    }
    return d;
}

sift3d_detector *sift3d_make_detector(void) { return sift3d_make_detector_on_device(-1); }

void sift3d_free_detector(sift3d_detector *d) {
    if (d == NULL)
        return;
    sift3d_cuda_host_free(d->scratch);
    sift3d_cuda_multi_destroy(d->multi);
    sift3d_cuda_ctx_destroy(d->ctx);
    free(d);
}

void *sift3d_detector_cuda_ctx(sift3d_detector *d) { return d ? d->ctx : NULL; }

int sift3d_detector_set_peak_thresh(sift3d_detector *const d, const double peak_thresh) {
    if (peak_thresh <= 0.0 || peak_thresh > 1) {
        ERR("sift3d_detector peak_thresh must be in the interval (0, 1]. Provided: %f \n",
            peak_thresh);
        return SIFT3D_FAILURE;
    }
    d->peak_thresh = peak_thresh;
    return SIFT3D_SUCCESS;
}

int sift3d_detector_set_corner_thresh(sift3d_detector *const d, const double corner_thresh) {
    if (corner_thresh < 0.0 || corner_thresh > 1.0) {
        ERR("sift3d_detector corner_thresh must be in the interval [0, 1]. Provided: %f \n",
            corner_thresh);
        return SIFT3D_FAILURE;
    }
    d->corner_thresh = corner_thresh;
    return SIFT3D_SUCCESS;
}

/* Parameter changes that reshape the pyramid or its filters take effect through a
 * re-plan. The reference re-derives them immediately when an image is present
 * (sift.c:467-474, 490-495); doing it lazily at the next detect is equivalent because
 * nothing can observe the pyramid in between except extract, which needs a detect. */
int sift3d_detector_set_num_kp_levels(sift3d_detector *const d, const unsigned int num_kp_levels) {
    if (num_kp_levels < 1 || num_kp_levels + 2 > SIFT3D_CUDA_MAX_FILTERS) {
        ERR("sift3d_detector num_kp_levels must be in [1, %d]. Provided: %u \n",
            SIFT3D_CUDA_MAX_FILTERS - 2, num_kp_levels);
        return SIFT3D_FAILURE;
    }
    d->num_kp_levels = (int) num_kp_levels;
    d->plan_valid = 0;
    d->have_gpyr = 0;
    return SIFT3D_SUCCESS;
}

static int set_scales(sift3d_detector *const d, const double sigma0, const double sigma_n) {
    /* set_scales_Pyramid only validates when the pyramid has levels, i.e. once an image
     * has been seen (imutil.c:1574-1588); before that the values are stored as they are. */
    if (d->have_dims && sigma0 * pow(2.0, -1.0 / d->num_kp_levels) < sigma_n) {
        ERR("set_scales_Pyramid: sigma_n too large for these settings. Max allowed: %f \n",
            sigma0 * pow(2.0, -1.0 / d->num_kp_levels) - DBL_EPSILON);
        return SIFT3D_FAILURE;
    }
    d->sigma0 = sigma0;
    d->sigma_n = sigma_n;
    d->plan_valid = 0;
    d->have_gpyr = 0;
    return SIFT3D_SUCCESS;
}

int sift3d_detector_set_sigma_n(sift3d_detector *const d, const double sigma_n) {
    if (sigma_n < 0.0) {
        ERR("sift3d_detector sigma_n must be nonnegative. Provided: %f \n", sigma_n);
        return SIFT3D_FAILURE;
    }
    return set_scales(d, d->sigma0, sigma_n);
}

int sift3d_detector_set_sigma0(sift3d_detector *const d, const double sigma0) {
    if (sigma0 < 0.0) {
        ERR("sift3d_detector sigma0 must be nonnegative. Provided: %f \n", sigma0);
        return SIFT3D_FAILURE;
    }
    return set_scales(d, sigma0, d->sigma_n);
}

/* ---- keypoint store ---------------------------------------------------------------- */

sift3d_keypoint_store *sift3d_make_keypoint_store(void) {
    return (sift3d_keypoint_store *) calloc(1, sizeof(sift3d_keypoint_store));
}

void sift3d_free_keypoint_store(sift3d_keypoint_store *kp) {
    if (kp == NULL)
        return;
    free(kp->buf);
    free(kp);
}

static int kp_reserve(sift3d_keypoint_store *kp, size_t num) {
    if (num > kp->cap) {
        /* grow in slabs of 500 like SIFT3D_RESIZE_SLAB (immacros.h:199-222) */
        const size_t cap = (num + 499) / 500 * 500;
        keypoint_rec *p = (keypoint_rec *) realloc(kp->buf, cap * sizeof(keypoint_rec));
        if (p == NULL)
            return SIFT3D_FAILURE;
        kp->buf = p;
        kp->cap = cap;
    }
    return SIFT3D_SUCCESS;
}

int sift3d_keypoint_store_count(const sift3d_keypoint_store *kp) { return kp ? (int) kp->num : 0; }

int sift3d_keypoint_store_get(const sift3d_keypoint_store *kp, int index, int *o, int *s,
                              double xyz[3], double *sd, float *strength, float R[9]) {
    const keypoint_rec *k;
    if (kp == NULL || index < 0 || (size_t) index >= kp->num)
        return SIFT3D_FAILURE;
    k = kp->buf + index;
    if (o) *o = k->o;
    if (s) *s = k->s;
    if (xyz) { xyz[0] = k->xd; xyz[1] = k->yd; xyz[2] = k->zd; }
    if (sd) *sd = k->sd;
    if (strength) *strength = k->strength;
    if (R) memcpy(R, k->R, sizeof(k->R));
    return SIFT3D_SUCCESS;
}

int sift3d_keypoint_store_to_mat_rm(const sift3d_keypoint_store *const kp, sift3d_mat_rm *const mat) {
    const int num = (int) kp->num;
    int i;

    if (mat_resize(mat, num, IM_NDIMS, SIFT3D_DOUBLE))
        return SIFT3D_FAILURE;
    for (i = 0; i < num; i++) {
        const keypoint_rec *const key = kp->buf + i;
        const double coord_factor = ldexp(1.0, key->o);
        double *row = (double *) mat->data + (size_t) i * IM_NDIMS;
        row[0] = coord_factor * key->xd;
        row[1] = coord_factor * key->yd;
        row[2] = coord_factor * key->zd;
    }
    return SIFT3D_SUCCESS;
}

int sift3d_keypoint_store_save(const char *path, const sift3d_keypoint_store *const kp) {
    /* columns: strength, x, y, z, o, sd, R row-major (sift.c:1746-1788) */
    sift3d_mat_rm mat;
    const int num_cols = 6 + IM_NDIMS * IM_NDIMS;
    int i, j, ret;

    memset(&mat, 0, sizeof(mat));
    if (mat_resize(&mat, (int) kp->num, num_cols, SIFT3D_DOUBLE))
        return SIFT3D_FAILURE;
    for (i = 0; i < (int) kp->num; i++) {
        const keypoint_rec *const key = kp->buf + i;
        double *row = (double *) mat.data + (size_t) i * num_cols;
        row[0] = key->strength;
        row[1] = key->xd;
        row[2] = key->yd;
        row[3] = key->zd;
        row[4] = key->o;
        row[5] = key->sd;
        for (j = 0; j < IM_NDIMS * IM_NDIMS; j++)
            row[6 + j] = (double) key->R[j];
    }
    ret = mat_write(path, &mat);
    free(mat.data);
    return ret;
}

/* keypoint_strength_cmp (sift.c:1832-1837) */
static int keypoint_strength_cmp(const void *p1, const void *p2) {
    const keypoint_rec *kp1 = (const keypoint_rec *) p1;
    const keypoint_rec *kp2 = (const keypoint_rec *) p2;
    return (kp1->strength < kp2->strength) ? 1 : -1;
}

void sift3d_keypoint_store_sort_by_strength(sift3d_keypoint_store *const store, int limit) {
    qsort(store->buf, store->num, sizeof(keypoint_rec), keypoint_strength_cmp);
    /* `num > limit` compares size_t with int in the reference (sift.c:1897): a negative
     * limit converts to a huge unsigned value and truncates nothing. */
    if (limit != 0 && store->num > (size_t) limit)
        store->num = (size_t) limit;
}

/* ---- descriptor store -------------------------------------------------------------- */

sift3d_descriptor_store *sift3d_make_descriptor_store(void) {
    return (sift3d_descriptor_store *) calloc(1, sizeof(sift3d_descriptor_store));
}

void sift3d_free_descriptor_store(sift3d_descriptor_store *ds) {
    if (ds == NULL)
        return;
    sift3d_cuda_host_free(ds->desc);
    free(ds->xyzs);
    free(ds);
}

int sift3d_descriptor_store_count(const sift3d_descriptor_store *ds) { return ds ? (int) ds->num : 0; }

int sift3d_descriptor_store_to_mat_rm(const sift3d_descriptor_store *const store,
                                      sift3d_mat_rm *const mat) {
    const int num_rows = (int) store->num;
    const int num_cols = IM_NDIMS + DESC_NUMEL;
    int i;

    if (num_rows < 1) {
        printf("SIFT3D_Descriptor_store_to_Mat_rm: invalid number of descriptors: %d \n", num_rows);
        return SIFT3D_FAILURE;
    }
    if (mat_resize(mat, num_rows, num_cols, SIFT3D_FLOAT))
        return SIFT3D_FAILURE;
    for (i = 0; i < num_rows; i++) {
        float *row = (float *) mat->data + (size_t) i * num_cols;
        row[0] = (float) store->xyzs[4 * (size_t) i + 0];
        row[1] = (float) store->xyzs[4 * (size_t) i + 1];
        row[2] = (float) store->xyzs[4 * (size_t) i + 2];
        /* column = hist_idx * 12 + bin + 3 (DESC_MAT_GET_COL, sift.c:136-137): the
         * device already stores elements in that order */
        memcpy(row + IM_NDIMS, store->desc + (size_t) i * DESC_NUMEL, DESC_NUMEL * sizeof(float));
    }
    return SIFT3D_SUCCESS;
}

int sift3d_descriptor_store_save(const char *path, const sift3d_descriptor_store *const desc) {
    sift3d_mat_rm mat;
    int ret;

    memset(&mat, 0, sizeof(mat));
    mat.type = SIFT3D_FLOAT;
    if (sift3d_descriptor_store_to_mat_rm(desc, &mat)) {
        free(mat.data);
        return SIFT3D_FAILURE;
    }
    ret = mat_write(path, &mat);
    free(mat.data);
    return ret;
}

/* ---- the hot path -------------------------------------------------------------------- */

int sift3d_detect_keypoints(sift3d_detector *const d, const sift3d_image *const im,
                            sift3d_keypoint_store *const kp) {
    double units[3];
    int num_cand = 0, i, n;
    size_t num;

    if (im->nc != 1) {
        ERR("SIFT3D_detect_keypoints: invalid number of image channels: %d -- only "
            "single-channel images are supported \n", im->nc);
        return SIFT3D_FAILURE;
    }
    if (im->data == NULL) /* im_copy_data, imutil.c:653-654 */
        return SIFT3D_FAILURE;
    units[0] = im->ux;
    units[1] = im->uy;
    units[2] = im->uz;

    /* set_im_SIFT3D: re-plan only when the dims changed (sift.c:651-656) or a parameter
     * that shapes the pyramid did */
    if (!d->have_dims || d->nx != im->nx || d->ny != im->ny || d->nz != im->nz || !d->plan_valid) {
        d->nx = im->nx;
        d->ny = im->ny;
        d->nz = im->nz;
        d->have_dims = 1;
        d->plan_valid = 0;
        d->have_gpyr = 0;
        if (detector_plan(d, units))
            return SIFT3D_FAILURE;
        d->plan_valid = 1;
    }

    {
        const int dev = sift3d_cuda_ctx_device(d->ctx);
        const int resident = im->dev != NULL && im->dev_id == dev;
        const float *src = resident ? (const float *) im->dev : im->data;
        int on_device = resident;
        if (im->up_pending && sift3d_cuda_upload_device(im->up) == dev) {
            /* a prefetched copy: the stream waits for the transfer, the host does not */
            if (sift3d_cuda_upload_wait(d->ctx, im->up))
                return SIFT3D_FAILURE;
            src = sift3d_cuda_upload_ptr(im->up);
            on_device = 1;
            ((sift3d_image *) im)->up_pending = 0; /* internal transfer state, not image data */
        }
        if (sift3d_cuda_detect(d->ctx, src, on_device, units, d->peak_thresh, d->corner_thresh,
                               &num_cand))
            return SIFT3D_FAILURE;
    }

    d->have_gpyr = 1;

    /* detect_extrema records the dims of the first octave (sift.c:756-759) */
    kp->nx = im->nx;
    kp->ny = im->ny;
    kp->nz = im->nz;

    /* Survivors are compacted in candidate order (sift.c:1148-1166). copy_Keypoint does
     * not copy `strength` (sift.c:372-384), so survivor j ends up with the strength that
     * candidate j had -- reproduced here because sort_by_strength() and the keypoint CSV
     * expose it. */
    if (kp_reserve(kp, (size_t) num_cand))
        return SIFT3D_FAILURE;
    num = 0;
    {
        const sift3d_cuda_keypoint *rec = NULL;
        const sift3d_cuda_candidate_flag *flg = NULL;
        if (sift3d_cuda_candidates_view(d->ctx, 0, &rec) >= 0 &&
            sift3d_cuda_candidates_flags(d->ctx, 0, &flg) >= 0) {
            /* The records are already in host memory (written there by the orientation
             * kernel), one region per octave in scan order. The 8-byte flags say which
             * candidates survived and carry every candidate's strength, so only the
             * survivors' 88-byte records are read. `sflg`/`sj` walk the flags a second time,
             * `num` entries behind, for the strength of candidate number `num`. */
            const sift3d_cuda_candidate_flag *sflg = flg;
            int o = 0, so = 0, j, sj = 0, cnt, scnt, seen = 0;
            cnt = scnt = sift3d_cuda_candidates_flags(d->ctx, 0, &flg);
            while (seen < num_cand) {
                while (cnt == 0) {
                    cnt = sift3d_cuda_candidates_flags(d->ctx, ++o, &flg);
                    if (cnt < 0 || sift3d_cuda_candidates_view(d->ctx, o, &rec) != cnt)
                        return SIFT3D_FAILURE;
                }
                for (j = 0; j < cnt; j++) {
                    const sift3d_cuda_keypoint *c;
                    keypoint_rec *dst;
                    if (!flg[j].valid)
                        continue;
                    while (sj >= scnt) {
                        scnt = sift3d_cuda_candidates_flags(d->ctx, ++so, &sflg);
                        sj = 0;
                    }
                    c = rec + j;
                    dst = kp->buf + num;
                    dst->xd = c->x;
                    dst->yd = c->y;
                    dst->zd = c->z;
                    dst->sd = c->sd;
                    dst->o = c->o;
                    dst->s = c->s;
                    memcpy(dst->R, c->R, sizeof(dst->R));
                    dst->strength = sflg[sj++].strength;
                    num++;
                }
                seen += cnt;
                cnt = 0;
            }
            kp->num = num;
            return SIFT3D_SUCCESS;
        }
    }

    if ((size_t) num_cand > d->scratch_cap) {
        /* page-locked, so the candidate download runs at PCIe speed */
        sift3d_cuda_host_free(d->scratch);
        d->scratch_cap = (size_t) num_cand * 2 + 1024;
        d->scratch = (sift3d_cuda_keypoint *)
            sift3d_cuda_host_alloc(d->scratch_cap * sizeof(sift3d_cuda_keypoint));
        if (d->scratch == NULL) {
            d->scratch_cap = 0;
            return SIFT3D_FAILURE;
        }
    }
    n = sift3d_cuda_fetch_candidates(d->ctx, d->scratch, num_cand);
    if (n != num_cand)
        return SIFT3D_FAILURE;
    for (i = 0; i < num_cand; i++) {
        const sift3d_cuda_keypoint *c = d->scratch + i;
        keypoint_rec *dst;
        if (!c->valid)
            continue;
        dst = kp->buf + num;
        dst->xd = c->x;
        dst->yd = c->y;
        dst->zd = c->z;
        dst->sd = c->sd;
        dst->o = c->o;
        dst->s = c->s;
        memcpy(dst->R, c->R, sizeof(dst->R));
        dst->strength = d->scratch[num].strength;
        num++;
    }
    kp->num = num;
    return SIFT3D_SUCCESS;
}

/* verify_keys (sift.c:1171-1212) */
static int verify_keys(const sift3d_keypoint_store *const kp, const sift3d_detector *d) {
    const int num = (int) kp->num;
    int i;

    if (num < 1) {
        ERR("verify_keys: invalid number of keypoints: %d \n", num);
        return SIFT3D_FAILURE;
    }
    for (i = 0; i < num; i++) {
        const keypoint_rec *key = kp->buf + i;
        const double octave_factor = ldexp(1.0, key->o);
        if (key->xd < 0 || key->yd < 0 || key->zd < 0 ||
            key->xd * octave_factor >= (double) d->nx ||
            key->yd * octave_factor >= (double) d->ny ||
            key->zd * octave_factor >= (double) d->nz) {
            ERR("verify_keys: keypoint %d (%f, %f, %f) octave %d exceeds image dimensions "
                "(%d, %d, %d) \n", i, key->xd, key->yd, key->zd, key->o, d->nx, d->ny, d->nz);
            return SIFT3D_FAILURE;
        }
        if (key->sd <= 0) {
            ERR("verify_keys: keypoint %d has invalid scale %f \n", i, key->sd);
            return SIFT3D_FAILURE;
        }
    }
    return SIFT3D_SUCCESS;
}

int sift3d_extract_descriptors(sift3d_detector *const d, const sift3d_keypoint_store *const kp,
                               sift3d_descriptor_store *const desc) {
    sift3d_cuda_keypoint *keys;
    float *dbuf;
    double *xbuf;
    int i, num;

    if (verify_keys(kp, d))
        return SIFT3D_FAILURE;
    if (!d->have_gpyr) {
        ERR("SIFT3D_extract_descriptors: no Gaussian pyramid is available. Make sure "
            "SIFT3D_detect_keypoints was called prior to calling this function. \n");
        return SIFT3D_FAILURE;
    }
    num = (int) kp->num;
    keys = (sift3d_cuda_keypoint *) malloc((size_t) num * sizeof(*keys));
    if ((size_t) num > desc->cap) {
        /* page-locked, so the descriptor download runs at PCIe speed */
        sift3d_cuda_host_free(desc->desc);
        desc->cap = (size_t) num + (size_t) num / 2 + 64;
        desc->desc = (float *) sift3d_cuda_host_alloc(desc->cap * DESC_NUMEL * sizeof(float));
        if (desc->desc == NULL)
            desc->cap = 0;
    }
    dbuf = desc->desc;
    xbuf = (double *) realloc(desc->xyzs, (size_t) num * 4 * sizeof(double));
    if (xbuf != NULL)
        desc->xyzs = xbuf;
    if (keys == NULL || dbuf == NULL || xbuf == NULL) {
        free(keys);
        desc->num = 0;
        return SIFT3D_FAILURE;
    }
    for (i = 0; i < num; i++) {
        const keypoint_rec *key = kp->buf + i;
        keys[i].o = key->o;
        keys[i].s = key->s;
        keys[i].x = key->xd;
        keys[i].y = key->yd;
        keys[i].z = key->zd;
        keys[i].sd = key->sd;
        keys[i].strength = key->strength;
        memcpy(keys[i].R, key->R, sizeof(key->R));
        keys[i].valid = 1;
    }
    /* metadata = dims of the first pyramid level (sift.c:1572-1575) */
    desc->nx = d->nx;
    desc->ny = d->ny;
    desc->nz = d->nz;
    if (sift3d_cuda_describe(d->ctx, keys, num, desc->desc, 0)) {
        free(keys);
        desc->num = 0;
        return SIFT3D_FAILURE;
    }
    for (i = 0; i < num; i++) {
        /* descriptor location in base-octave coordinates (sift.c:1528-1533) */
        const keypoint_rec *key = kp->buf + i;
        const double coord_factor = ldexp(1.0, key->o);
        desc->xyzs[4 * (size_t) i + 0] = key->xd * coord_factor;
        desc->xyzs[4 * (size_t) i + 1] = key->yd * coord_factor;
        desc->xyzs[4 * (size_t) i + 2] = key->zd * coord_factor;
        desc->xyzs[4 * (size_t) i + 3] = key->sd;
    }
    desc->num = (size_t) num;
    free(keys);
    return SIFT3D_SUCCESS;
}

/* ---- batch: many volumes over several detectors (devices) ------------------------------
 * Extension; the per-volume work is exactly sift3d_detect_keypoints
 * + sift3d_extract_descriptors (sift.h:92-111), so every result equals the one-volume calls.
 * Volume i goes to detector i % num_det; every detector is driven by a host thread of its own
 * (one detector = one device + one stream), and the upload of a detector's next volume
 * (sift3d_image_prefetch) is started between detect and extract of the current one, where it
 * hides under the descriptor kernel. A volume without keypoints gets an empty descriptor
 * store (the one-volume extract call fails on an empty store, sift.c:1178-1182). */
#include <pthread.h>

typedef struct {
    sift3d_detector *det;
    sift3d_image *const *ims;
    sift3d_keypoint_store *const *kps;
    sift3d_descriptor_store *const *descs;
    int first, stride, num;
    int status;
} batch_job;

static void *batch_worker(void *arg) {
    batch_job *const j = (batch_job *) arg;
    const int dev = sift3d_cuda_ctx_device(j->det->ctx);
    int i;
    j->status = SIFT3D_SUCCESS;
    if (j->first < j->num)
        sift3d_image_prefetch(j->ims[j->first], dev);
    for (i = j->first; i < j->num; i += j->stride) {
        if (sift3d_detect_keypoints(j->det, j->ims[i], j->kps[i])) {
            j->status = SIFT3D_FAILURE;
            break;
        }
        if (i + j->stride < j->num)
            sift3d_image_prefetch(j->ims[i + j->stride], dev);
        if (j->descs == NULL || j->descs[i] == NULL)
            continue;
        if (j->kps[i]->num == 0) {
            j->descs[i]->num = 0;
            continue;
        }
        if (sift3d_extract_descriptors(j->det, j->kps[i], j->descs[i])) {
            j->status = SIFT3D_FAILURE;
            break;
        }
    }
    return NULL;
}

int sift3d_detect_describe_batch(sift3d_detector *const *dets, int num_det,
                                 sift3d_image *const *ims, sift3d_keypoint_store *const *kps,
                                 sift3d_descriptor_store *const *descs, int num) {
    pthread_t *th;
    batch_job *jobs;
    int d, ret = SIFT3D_SUCCESS;
    if (dets == NULL || ims == NULL || kps == NULL || num_det < 1 || num < 0)
        return SIFT3D_FAILURE;
    for (d = 0; d < num_det; d++)
        if (dets[d] == NULL || dets[d]->ctx == NULL)
            return SIFT3D_FAILURE;
    for (d = 0; d < num; d++)
        if (ims[d] == NULL || kps[d] == NULL)
            return SIFT3D_FAILURE;
    if (num == 0)
        return SIFT3D_SUCCESS;
    if (num_det > num)
        num_det = num;
    th = (pthread_t *) calloc((size_t) num_det, sizeof(*th));
    jobs = (batch_job *) calloc((size_t) num_det, sizeof(*jobs));
    if (th == NULL || jobs == NULL) {
        free(th);
        free(jobs);
        return SIFT3D_FAILURE;
    }
    for (d = 0; d < num_det; d++) {
        jobs[d].det = dets[d];
        jobs[d].ims = ims;
        jobs[d].kps = kps;
        jobs[d].descs = descs;
        jobs[d].first = d;
        jobs[d].stride = num_det;
        jobs[d].num = num;
        jobs[d].status = SIFT3D_FAILURE;
    }
    /* detector 0 runs on the calling thread */
    for (d = 1; d < num_det; d++)
        if (pthread_create(&th[d], NULL, batch_worker, &jobs[d])) {
            jobs[d].stride = 0; /* not started */
            ret = SIFT3D_FAILURE;
        }
    batch_worker(&jobs[0]);
    for (d = 1; d < num_det; d++)
        if (jobs[d].stride != 0)
            pthread_join(th[d], NULL);
    for (d = 0; d < num_det; d++)
        if (jobs[d].status != SIFT3D_SUCCESS)
            ret = SIFT3D_FAILURE;
    free(th);
    free(jobs);
    return ret;
}

/* ---- upstream-cased aliases (see include/sift3d.h) -------------------------------------- */
int SIFT3D_detect_keypoints(SIFT3D *d, const Image *im, Keypoint_store *kp) {
    return sift3d_detect_keypoints(d, im, kp);
}

int SIFT3D_extract_descriptors(SIFT3D *d, const Keypoint_store *kp, SIFT3D_Descriptor_store *desc) {
    return sift3d_extract_descriptors(d, kp, desc);
}

/* ---- one tall volume over several devices (z-slabs) --------------------------------------
 * Extension: the same path as sift3d_detect_keypoints +
 * sift3d_extract_descriptors (sift.h:92-111), decomposed over z by libsift3d_cuda.so's
 * sift3d_cuda_multi (multi.cu). This side computes the layout, designs the per-slab plans with
 * the detector's parameters and merges the per-device results in the reference's scan order
 * (o, s, z, y, x) = (o, s, device, local order), compacting the survivors like
 * assign_orientations (sift.c:1148-1166), strength quirk included. */
#define DESC_SIG_FCTR 7.071067812 /* sift.c:43 */
#define DESC_RAD_FCTR 2.0         /* sift.c:44 */

/* Layout of `nz` planes over `num` devices: owned ranges on multiples of 2^K, each with `halo`
 * planes on its cut faces. Returns the halo, or -1 if the volume cannot be split that way. */
int sift3d_sharded_layout(int nz, int num, int K, int window_planes, int max_half_width, double uz,
                          int *g0, int *g1, int *a0, int *a1) {
    const int align = 1 << K;
    const int nblk = nz / align;
    int reach = (int) ceil((double) max_half_width / uz) + 2, halo, r;
    if (reach < window_planes)
        reach = window_planes;
    halo = (reach << (K - 1)) + align - 1;
    halo -= halo % align;
    if (num < 1 || K < 1 || nblk < num)
        return -1;
    for (r = 0; r < num; r++) {
        g0[r] = align * (int) (((long) r * nblk) / num);
        g1[r] = r + 1 < num ? align * (int) (((long) (r + 1) * nblk) / num) : nz;
        a0[r] = g0[r] - halo > 0 ? g0[r] - halo : 0;
        a1[r] = g1[r] + halo < nz ? g1[r] + halo : nz;
    }
    return halo;
}

int sift3d_detect_describe_sharded(sift3d_detector *const d, const sift3d_image *const im,
                                   const int *devices, int num_devices, int sharded_octaves,
                                   sift3d_keypoint_store *const kp,
                                   sift3d_descriptor_store *const desc) {
    sift3d_cuda_multi_desc M;
    double key[32];
    int total, K, tail_octaves, window, hwmax, r, o, s, i, part;
    size_t ncand = 0, num = 0;
    int idx[SIFT3D_CUDA_MULTI_MAX_DEVICES + 1];
    const sift3d_cuda_keypoint *cands[SIFT3D_CUDA_MULTI_MAX_DEVICES + 1];
    int *rows[SIFT3D_CUDA_MULTI_MAX_DEVICES + 1];
    int counts[SIFT3D_CUDA_MULTI_MAX_DEVICES + 1], vidx[SIFT3D_CUDA_MULTI_MAX_DEVICES + 1];
    float *strengths = NULL;
    int ok;

    if (d == NULL || im == NULL || devices == NULL || kp == NULL || num_devices < 1 ||
        num_devices > SIFT3D_CUDA_MULTI_MAX_DEVICES)
        return SIFT3D_FAILURE;
    if (im->nc != 1 || im->data == NULL) {
        ERR("sift3d_detect_describe_sharded: single-channel images only \n");
        return SIFT3D_FAILURE;
    }
    {
        int min_dim = im->nx < im->ny ? im->nx : im->ny;
        min_dim = min_dim < im->nz ? min_dim : im->nz;
        total = (int) log2((double) min_dim) - 3 + 1; /* sift.c:441-452 */
    }
    if (total < 1) {
        ERR("resize_SIFT3D: input image is too small: must have at least 8 voxels in each "
            "dimension \n");
        return SIFT3D_FAILURE;
    }
    K = sharded_octaves < 1 ? 1 : (sharded_octaves > total ? total : sharded_octaves);
    tail_octaves = total - K;

    /* (re)build the multi-device object when anything it depends on changed */
    memset(key, 0, sizeof(key));
    key[0] = im->nx; key[1] = im->ny; key[2] = im->nz;
    key[3] = im->ux; key[4] = im->uy; key[5] = im->uz;
    key[6] = d->sigma_n; key[7] = d->sigma0; key[8] = d->num_kp_levels;
    key[9] = K; key[10] = num_devices;
    for (r = 0; r < num_devices; r++)
        key[11 + r] = devices[r];
    if (d->multi == NULL || memcmp(key, d->multi_key, sizeof(key)) != 0) {
        sift3d_cuda_plan_desc *plans, *tail_plan = NULL;
        const double units[3] = {im->ux, im->uy, im->uz};
        int ok = 1;
        sift3d_cuda_multi_destroy(d->multi);
        d->multi = NULL;
        plans = (sift3d_cuda_plan_desc *) calloc((size_t) num_devices + 1, sizeof(*plans));
        if (plans == NULL)
            return SIFT3D_FAILURE;
        /* filters do not depend on the geometry: design once to learn the blur's reach */
        if (fill_plan(d, im->nx, im->ny, im->nz, units, 0, 1, &plans[0])) {
            free(plans);
            return SIFT3D_FAILURE;
        }
        hwmax = plans[0].first.width / 2;
        for (i = 0; i < d->num_kp_levels + 2; i++)
            if (plans[0].octave[i].width / 2 > hwmax)
                hwmax = plans[0].octave[i].width / 2;
        /* widest window of the path in planes of ITS octave: the descriptor window radius
         * desc_rad_fctr * desc_sig_fctr * sd / unit of the top keypoint level (sift.c:1453-1461;
         * it dominates the orientation window, sift.c:936), + 1 for the central differences,
         * + 2 for the ceil / floor of the window bounds */
        window = (int) ceil(DESC_RAD_FCTR * DESC_SIG_FCTR * d->sigma0 *
                            pow(2.0, (double) (d->num_kp_levels - 1) / d->num_kp_levels) / im->uz) + 3;
        memset(&M, 0, sizeof(M));
        M.num_devices = num_devices;
        for (r = 0; r < num_devices; r++)
            M.devices[r] = devices[r];
        M.nx = im->nx; M.ny = im->ny; M.nz = im->nz;
        memcpy(M.units, units, sizeof(units));
        M.sharded_octaves = K;
        for (o = 0; o < SIFT3D_CUDA_MAX_OCTAVES; o++)
            M.window_planes[o] = window;
        if (sift3d_sharded_layout(im->nz, num_devices, K, window, hwmax, im->uz, M.g0, M.g1, M.a0, M.a1) < 0) {
            ERR("sift3d_detect_describe_sharded: %d planes cannot be split into %d slabs aligned to %d \n",
                im->nz, num_devices, 1 << K);
            free(plans);
            return SIFT3D_FAILURE;
        }
        for (r = 0; r < num_devices && ok; r++)
            ok = fill_plan(d, im->nx, im->ny, M.a1[r] - M.a0[r], units, 0, K + (tail_octaves ? 1 : 0),
                           &plans[r]) == SIFT3D_SUCCESS;
        if (ok && tail_octaves) {
            const double tu[3] = {ldexp(units[0], K), ldexp(units[1], K), ldexp(units[2], K)};
            tail_plan = &plans[num_devices];
            ok = fill_plan(d, im->nx >> K, im->ny >> K, im->nz >> K, tu, K, tail_octaves, tail_plan) ==
                 SIFT3D_SUCCESS;
        }
        if (ok)
            d->multi = sift3d_cuda_multi_create(&M, plans, tail_plan);
        free(plans);
        if (d->multi == NULL)
            return SIFT3D_FAILURE;
        memcpy(d->multi_key, key, sizeof(key));
    }

    if (sift3d_cuda_multi_detect(d->multi, im->data, d->peak_thresh, d->corner_thresh))
        return SIFT3D_FAILURE;

    /* ---- merge: (o, s, device, local order); the unsharded octaves follow */
    for (part = 0; part <= num_devices; part++) {
        counts[part] = sift3d_cuda_multi_results(d->multi, part, &cands[part]);
        if (counts[part] < 0)
            return SIFT3D_FAILURE;
        ncand += (size_t) counts[part];
        idx[part] = vidx[part] = 0;
        rows[part] = NULL;
    }
    kp->nx = im->nx;
    kp->ny = im->ny;
    kp->nz = im->nz;
    if (kp_reserve(kp, ncand))
        return SIFT3D_FAILURE;
    strengths = (float *) malloc((ncand + 1) * sizeof(float));
    ok = strengths != NULL;
    for (part = 0; part <= num_devices && ok; part++) {
        /* row of the j-th valid candidate of the part in the merged survivor list */
        rows[part] = (int *) malloc(((size_t) counts[part] + 1) * sizeof(int));
        ok = rows[part] != NULL;
    }
    if (ok) {
        size_t seen = 0;
#define TAKE(PART)                                                                              \
    do {                                                                                        \
        const sift3d_cuda_keypoint *c = cands[PART] + idx[PART]++;                              \
        strengths[seen++] = c->strength;                                                        \
        if (c->valid) {                                                                         \
            keypoint_rec *dst = kp->buf + num;                                                  \
            dst->xd = c->x; dst->yd = c->y; dst->zd = c->z; dst->sd = c->sd;                    \
            dst->o = c->o; dst->s = c->s;                                                       \
            memcpy(dst->R, c->R, sizeof(dst->R));                                               \
            rows[PART][vidx[PART]++] = (int) num;                                               \
            num++;                                                                              \
        }                                                                                       \
    } while (0)
        for (o = 0; o < K; o++)
            for (s = 0; s < d->num_kp_levels; s++)
                for (part = 0; part < num_devices; part++)
                    while (idx[part] < counts[part] && cands[part][idx[part]].o == o &&
                           cands[part][idx[part]].s == s)
                        TAKE(part);
        while (idx[num_devices] < counts[num_devices])
            TAKE(num_devices);
#undef TAKE
        if (seen != ncand) {
            ERR("sift3d_detect_describe_sharded: per-device candidate lists are not in scan order \n");
            ok = 0;
        }
    }
    if (ok) {
        /* copy_Keypoint does not copy `strength` (sift.c:372-384): survivor j carries candidate j's */
        for (i = 0; (size_t) i < num; i++)
            kp->buf[i].strength = strengths[i];
        kp->num = num;
    }
    if (ok && desc != NULL) {
        desc->num = 0;
        desc->nx = im->nx;
        desc->ny = im->ny;
        desc->nz = im->nz;
        if (num > desc->cap) {
            /* page-locked: every device writes its rows straight into it */
            sift3d_cuda_host_free(desc->desc);
            desc->cap = num + num / 2 + 64;
            desc->desc = (float *) sift3d_cuda_host_alloc(desc->cap * DESC_NUMEL * sizeof(float));
            if (desc->desc == NULL)
                desc->cap = 0;
        }
        if (num > 0) {
            double *xb = (double *) realloc(desc->xyzs, num * 4 * sizeof(double));
            if (xb != NULL)
                desc->xyzs = xb;
            ok = xb != NULL && desc->desc != NULL &&
                 sift3d_cuda_multi_describe(d->multi, desc->desc, (const int *const *) rows) == 0;
            for (i = 0; ok && (size_t) i < num; i++) {
                /* descriptor location in base-octave coordinates (sift.c:1528-1533) */
                const keypoint_rec *key = kp->buf + i;
                const double f = ldexp(1.0, key->o);
                desc->xyzs[4 * (size_t) i + 0] = key->xd * f;
                desc->xyzs[4 * (size_t) i + 1] = key->yd * f;
                desc->xyzs[4 * (size_t) i + 2] = key->zd * f;
                desc->xyzs[4 * (size_t) i + 3] = key->sd;
            }
        }
        if (ok)
            desc->num = num;
    }
    for (part = 0; part <= num_devices; part++)
        free(rows[part]);
    free(strengths);
    d->have_gpyr = 0; /* the pyramid lives on the slabs: no later extract on this detector */
    return ok ? SIFT3D_SUCCESS : SIFT3D_FAILURE;
}

/* Wall-clock split of the last sharded run (milliseconds): total, slab detects, unsharded
 * octaves' detect, describe phase, slowest slab detect, slowest part's describe, and the MiB of
 * halo planes moved between devices. */
int sift3d_sharded_timings(sift3d_detector *d, float out[7]) {
    sift3d_cuda_multi_timings t;
    if (d == NULL || d->multi == NULL || out == NULL || sift3d_cuda_multi_get_timings(d->multi, &t))
        return SIFT3D_FAILURE;
    out[0] = t.total_ms; out[1] = t.slabs_ms; out[2] = t.tail_ms; out[3] = t.describe_ms;
    out[4] = t.max_detect_ms; out[5] = t.max_describe_ms; out[6] = t.exchanged_mbytes;
    return SIFT3D_SUCCESS;
}
