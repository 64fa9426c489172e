/* imtypes.h -- source-compatibility shim: programs written against the reference's
 * <imtypes.h> (fatimp/SIFT3D sift3d/imtypes.h) compile unchanged with -I<repo>/include/sift3d.
 * All declarations live in ../sift3d.h. */
#ifndef SIFT3D_B200_COMPAT_IMTYPES_H
#define SIFT3D_B200_COMPAT_IMTYPES_H
#include "../sift3d.h"
#endif
