/*
 * nbody_constants.h -- the reference's compile-time parameters
 * (TestProject/constants.h:11-26 and, under NBODY_HEADLESS_CONSTANTS,
 * TestProject/Sim-Without-OpenGL-Integration/constants.h:3-15) kept under the
 * same macro names so that code written against the reference keeps compiling.
 * In this engine they are only DEFAULTS: every value is a run-time field of
 * nbody_params (include/nbody_b200.h).  Each macro can be overridden with -D.
 *
 * THREADS_PER_BLOCK and TILE_WIDTH_FACTOR are accepted and ignored: the block
 * shape belongs to the kernel variant chosen for the device (the reference's
 * 32x1 launch, kernel.cu:636-643, is what this engine replaces).
 */
#ifndef NBODY_CONSTANTS_H
#define NBODY_CONSTANTS_H

#ifndef THREADS_PER_BLOCK
#  define THREADS_PER_BLOCK 32
#endif
#ifndef TILE_WIDTH_FACTOR
#  define TILE_WIDTH_FACTOR 1
#endif

#ifdef NBODY_HEADLESS_CONSTANTS        /* double literals, as in the headless directory */
#  ifndef N_BODIES
#    define N_BODIES 5000
#  endif
#  ifndef MAX_W
#    define MAX_W 100000000.0
#  endif
#  ifndef EPS2
#    define EPS2 0.002
#  endif
#  ifndef DT
#    define DT 0.01
#  endif
#  define NBODY_BOX_ 100000.0
#  define NBODY_MINW_ 100000.0
#else                                   /* float literals, as in TestProject/constants.h */
#  ifndef N_BODIES
#    define N_BODIES 8192
#  endif
#  ifndef MAX_W
#    define MAX_W 1000000000.0f
#  endif
#  ifndef EPS2
#    define EPS2 0.002f
#  endif
#  ifndef DT
#    define DT 0.1f
#  endif
#  define NBODY_BOX_ 100000.0f
#  define NBODY_MINW_ 100000.0f
#endif

#ifndef MAX_X
#  define MAX_X NBODY_BOX_
#endif
#ifndef MAX_Y
#  define MAX_Y NBODY_BOX_
#endif
#ifndef MAX_Z
#  define MAX_Z NBODY_BOX_
#endif
#ifndef MIN_W
#  define MIN_W NBODY_MINW_
#endif

#endif /* NBODY_CONSTANTS_H */
