/* imutil.h -- source-compatibility shim: programs written against the reference's
 * <imutil.h> (fatimp/SIFT3D sift3d/imutil.h) compile unchanged with -I<repo>/include/sift3d.
 * All declarations live in ../sift3d.h. */
#ifndef SIFT3D_B200_COMPAT_IMUTIL_H
#define SIFT3D_B200_COMPAT_IMUTIL_H
#include "../sift3d.h"
#endif
