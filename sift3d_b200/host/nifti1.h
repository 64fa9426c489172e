/* nifti1.h -- minimal NIfTI-1 / Analyze 7.5 I/O for the host library (see nifti1.c). */
#ifndef SIFT3D_B200_NIFTI1_H
#define SIFT3D_B200_NIFTI1_H

#include <stddef.h>

typedef struct {
    int nx, ny, nz, nc;    /* nc = nt of a 4-D file, else 1 */
    double ux, uy, uz;     /* pixdim[1..3] */
} nifti1_info;

/* The voxel buffer comes from the caller (page-locked memory for images). */
typedef void *(*nifti1_alloc_fn)(size_t bytes, void *user);

/* 0 on success; -1 with a message on stderr. *data holds nx*ny*nz*nc floats laid out
 * channel-fastest: c + nc*(x + nx*(y + ny*z)). */
int nifti1_read(const char *path, nifti1_info *info, nifti1_alloc_fn alloc, void *user,
                float **data);

/* float32, slope 1, intercept 0, units in pixdim; .nii or .nii.gz. */
int nifti1_write(const char *path, const nifti1_info *info, const float *data);

#endif
