/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load the library built from this file.
 *
 * CPU restatement (plain C) of the rotated IoU + greedy rotated NMS that the
 * reference reaches through mmcv.ops (see rotgeo_ref_impl.inc for provenance
 * and the reference call sites).  Pinned against the reference's own golden
 * vectors in tests/test_oracle_golden.py (IoU).  NMS: the reference has no
 * test that pins any keep index -> "parity unpinned" for NMS, anchored on
 * Spec B only.
 *
 * Build: make -C oracle   (gcc -O3 -ffp-contract=off -fopenmp -shared)
 */
#include <math.h>
#include <stdlib.h>

#define T float
#define SFX _f32
#include "rotgeo_ref_impl.inc"
#undef T
#undef SFX

#define T double
#define SFX _f64
#include "rotgeo_ref_impl.inc"
#undef T
#undef SFX

#ifdef _OPENMP
#include <omp.h>
/* processors available to this process (affinity-aware); deliberately not omp_get_max_threads(): launchers
   such as torchrun export OMP_NUM_THREADS=1, and the callers pass an explicit thread count anyway */
int rotgeo_ref_max_threads(void) { return omp_get_num_procs(); }
#else
int rotgeo_ref_max_threads(void) { return 1; }
#endif
