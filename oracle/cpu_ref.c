/* cpu_ref.c -- C + OpenMP restatement of the reference's CPU path for the periodic explicit-RK step.
 *
 * TEST / BASELINE INFRASTRUCTURE ONLY: linked by nothing in the product.  Used by tests/ (checked
 * against the NumPy oracle oracle/ins_oracle.py, which is itself pinned to the reference's property and
 * known-answer tests) and by bench.py's `cpu_baseline` / `--impl reference` legs, where it stands for
 * "the reference's multithreaded CPU path" (KernelAbstractions CPU() kernels on Julia threads + FFTW):
 * the reference is Julia, and no Julia runtime exists in the build image or on the GPU box.
 *
 * It keeps the reference's LAUNCH STRUCTURE (one parallel loop per kernel / broadcast of
 * src/time_steppers/step_explicit_runge_kutta.jl:4-59, unfused) and is generous to the reference in two
 * ways: broadcasts are multithreaded here (Base broadcasts are single-threaded on the Julia CPU backend,
 * docs/src/manual/gpu.md:11-14) and there is no per-launch task-spawn overhead.
 * The FFTs go through a callback (scipy.fft / pocketfft with all workers; the reference uses FFTW.jl).
 *
 * Build: make -C oracle   ->  oracle/_build/libcpuref.so  (gcc -O3 -march=native -fopenmp -ffp-contract=off)
 */
#include <math.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define REAL float
#define SUF f32
#include "cpu_ref_impl.h"
#undef REAL
#undef SUF

#define REAL double
#define SUF f64
#include "cpu_ref_impl.h"
#undef REAL
#undef SUF

/* bench.py sets the thread count itself: torch.distributed.run exports OMP_NUM_THREADS=1 to its workers */
void cpuref_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int cpuref_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
