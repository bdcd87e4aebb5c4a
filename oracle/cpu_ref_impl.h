/* cpu_ref_impl.h -- body of oracle/cpu_ref.c, included once per precision (REAL, SUF).
 *
 * TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/cpu_ref.c).  One OpenMP-parallel loop per kernel or
 * broadcast of the reference's launch structure; arithmetic in REAL with the reference's operation
 * order (divisions by the spacing, /2 averages), compiled with -ffp-contract=off.
 * Arrays: reference layout, column-major (N1,..,ND[,D]), N = n + 2.  2-D: N3 = 1, no z terms.
 */
#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)

typedef struct {
  int D;
  int n[3], N[3];
  long s[3];   /* strides of the three axes */
  long sc;     /* component stride */
  int lo[3], hi[3]; /* interior index range per axis [lo, hi) */
  REAL h[3];
  REAL visc;
} FN(geo_);

static void FN(geo_init_)(FN(geo_) *g, int D, const int *n, const double *h, double Re) {
  g->D = D;
  for (int a = 0; a < 3; a++) {
    g->n[a] = a < D ? n[a] : 1;
    g->N[a] = a < D ? n[a] + 2 : 1;
    g->lo[a] = a < D ? 1 : 0;
    g->hi[a] = a < D ? n[a] + 1 : 1;
    g->h[a] = a < D ? (REAL)h[a] : (REAL)1;
  }
  g->s[0] = 1; g->s[1] = g->N[0]; g->s[2] = (long)g->N[0] * g->N[1];
  g->sc = (long)g->N[0] * g->N[1] * g->N[2];
  g->visc = (REAL)1 / (REAL)Re;
}

/* apply_bc_u! / apply_bc_p! (PeriodicBC): src/boundary_conditions.jl:276-288, 306-318.
 * nc components of `cs` words each (cs > 1: the SMatrix tensor field of the closure, AoS). */
static void FN(bc_fill_)(const FN(geo_) *g, REAL *f, int nc, int cs) {
  for (int b = 0; b < g->D; b++) {
    const int o1 = (b + 1) % 3, o2 = (b + 2) % 3;
    const long sb = g->s[b], s1 = g->s[o1], s2 = g->s[o2];
    const int Nb = g->N[b];
#pragma omp parallel for collapse(2) schedule(static)
    for (int c = 0; c < nc; c++)
      for (int j = 0; j < g->N[o2]; j++)
        for (int i = 0; i < g->N[o1]; i++) {
          REAL *p = f + (c * g->sc + i * s1 + j * s2) * (cs > 1 ? cs : 1);
          const long st = sb * (cs > 1 ? cs : 1);
          for (int w = 0; w < (cs > 1 ? cs : 1); w++) {
            p[w] = p[(Nb - 2) * st + w];
            p[(Nb - 1) * st + w] = p[st + w];
          }
        }
  }
}

static void FN(fill_)(REAL *f, long n, REAL v) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; i++) f[i] = v;
}

/* convection_diffusion_kernel!  src/operators.jl:647-690  (F accumulates; F was zero-filled) */
static void FN(convdiff_)(const FN(geo_) *g, REAL *F, const REAL *u) {
  const int D = g->D;
  const REAL two = (REAL)2, half = (REAL)0.5, visc = g->visc;
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = g->lo[2]; k < g->hi[2]; k++)
    for (int j = g->lo[1]; j < g->hi[1]; j++)
      for (int i = g->lo[0]; i < g->hi[0]; i++) {
        const long I = i + j * g->s[1] + k * g->s[2];
        for (int a = 0; a < D; a++) {
          const REAL *ua = u + a * g->sc;
          REAL f = F[a * g->sc + I];
          const REAL uI = ua[I];
          for (int b = 0; b < D; b++) {
            const REAL *ub = u + b * g->sc;
            const long eb = g->s[b], ea = g->s[a];
            const REAL hb = g->h[b];
            const REAL ua1 = (ua[I - eb] + uI) / two, ua2 = (uI + ua[I + eb]) / two;
            const REAL ub1 = half * ub[I - eb] + half * ub[I - eb + ea];
            const REAL ub2 = half * ub[I] + half * ub[I + ea];
            const REAL uu1 = ua1 * ub1, uu2 = ua2 * ub2;
            const REAL d1 = (uI - ua[I - eb]) / hb, d2 = (ua[I + eb] - uI) / hb;
            f += (visc * (d2 - d1) - (uu2 - uu1)) / hb;
          }
          F[a * g->sc + I] = f;
        }
      }
}

/* F .+= bodyforce (steady, stored field): src/operators.jl:861-868; also ku .+= s and the stage axpys */
static void FN(axpy_)(REAL *y, REAL a, const REAL *x, long n) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; i++) y[i] = y[i] + a * x[i];
}
static void FN(add_)(REAL *y, const REAL *x, long n) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; i++) y[i] = y[i] + x[i];
}
static void FN(copy_)(REAL *y, const REAL *x, long n) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; i++) y[i] = x[i];
}

/* dx(u, I, a, b): src/operators.jl:1023-1033 */
static inline REAL FN(ddx_)(const FN(geo_) *g, const REAL *u, long I, int a, int b) {
  const REAL *ua = u + a * g->sc;
  const long ea = g->s[a], eb = g->s[b];
  const REAL hb = g->h[b];
  if (a == b) return (ua[I] - ua[I - eb]) / hb;
  return ((ua[I + eb] - ua[I]) / hb + (ua[I - ea + eb] - ua[I - ea]) / hb + (ua[I] - ua[I - eb]) / hb +
          (ua[I - ea] - ua[I - ea - eb]) / hb) / (REAL)4;
}

/* smagtensor! kernel: src/operators.jl:1135-1151; sigma is AoS, D*D words per cell, column-major SMatrix */
static void FN(smagtensor_)(const FN(geo_) *g, REAL *sig, const REAL *u, REAL theta) {
  const int D = g->D, DD = D * D;
  REAL d2 = (REAL)0;
  for (int a = 0; a < D; a++) d2 = a == 0 ? g->h[a] * g->h[a] : d2 + g->h[a] * g->h[a];
  const REAL d = sizeof(REAL) == 4 ? (REAL)sqrtf((float)d2) : (REAL)sqrt((double)d2);  /* gridsize :1128 */
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = g->lo[2]; k < g->hi[2]; k++)
    for (int j = g->lo[1]; j < g->hi[1]; j++)
      for (int i = g->lo[0]; i < g->hi[0]; i++) {
        const long I = i + j * g->s[1] + k * g->s[2];
        REAL G[3][3], S[3][3];
        for (int a = 0; a < D; a++)
          for (int b = 0; b < D; b++) G[a][b] = FN(ddx_)(g, u, I, a, b);
        for (int a = 0; a < D; a++)
          for (int b = 0; b < D; b++) S[a][b] = (G[a][b] + G[b][a]) / (REAL)2;
        REAL ss = (REAL)0;
        int first = 1;
        for (int b = 0; b < D; b++)
          for (int a = 0; a < D; a++) {
            const REAL t = S[a][b] * S[a][b];
            ss = first ? t : ss + t;
            first = 0;
          }
        const REAL rt = sizeof(REAL) == 4 ? (REAL)sqrtf((float)((REAL)2 * ss)) : (REAL)sqrt((double)((REAL)2 * ss));
        const REAL eddy = theta * theta * (d * d) * rt;
        for (int b = 0; b < D; b++)
          for (int a = 0; a < D; a++) sig[I * DD + a + D * b] = (REAL)2 * eddy * S[a][b];
      }
}

/* divoftensor_kernel!  src/operators.jl:1205-1238 */
static void FN(divoftensor_)(const FN(geo_) *g, REAL *s, const REAL *sig) {
  const int D = g->D, DD = D * D;
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = g->lo[2]; k < g->hi[2]; k++)
    for (int j = g->lo[1]; j < g->hi[1]; j++)
      for (int i = g->lo[0]; i < g->hi[0]; i++) {
        const long I = i + j * g->s[1] + k * g->s[2];
        for (int a = 0; a < D; a++) {
          REAL acc = (REAL)0;
          const long ea = g->s[a];
          for (int b = 0; b < D; b++) {
            const long eb = g->s[b];
            const int c = a + D * b;
            REAL s1, s2;
            if (a == b) {
              s2 = sig[(I + eb) * DD + c];
              s1 = sig[I * DD + c];
            } else {
              s2 = (sig[I * DD + c] + sig[(I + eb) * DD + c] + sig[(I + ea + eb) * DD + c] + sig[(I + ea) * DD + c]) / (REAL)4;
              s1 = (sig[(I - eb) * DD + c] + sig[I * DD + c] + sig[(I + ea - eb) * DD + c] + sig[(I + ea) * DD + c]) / (REAL)4;
            }
            acc += (s2 - s1) / g->h[b];
          }
          s[a * g->sc + I] = acc;
        }
      }
}

/* divergence!  src/operators.jl:106-125 */
static void FN(divergence_)(const FN(geo_) *g, REAL *div, const REAL *u) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = g->lo[2]; k < g->hi[2]; k++)
    for (int j = g->lo[1]; j < g->hi[1]; j++)
      for (int i = g->lo[0]; i < g->hi[0]; i++) {
        const long I = i + j * g->s[1] + k * g->s[2];
        REAL d = (REAL)0;
        for (int a = 0; a < g->D; a++) d += (u[a * g->sc + I] - u[a * g->sc + I - g->s[a]]) / g->h[a];
        div[I] = d;
      }
}

/* scalewithvolume!  src/operators.jl:81-95 (whole array, volume formed first) */
static void FN(scale_)(REAL *p, long n, REAL v) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; i++) p[i] = p[i] * v;
}

/* copyto!(pI, view(p, Ip)) / copyto!(view(p, Ip), pI): src/pressure.jl:320, 347 */
static void FN(strip_)(const FN(geo_) *g, REAL *pI, REAL *p, int back) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = g->lo[2]; k < g->hi[2]; k++)
    for (int j = g->lo[1]; j < g->hi[1]; j++)
      for (int i = g->lo[0]; i < g->hi[0]; i++) {
        const long I = i + j * g->s[1] + k * g->s[2];
        const long J = (i - g->lo[0]) + (long)g->n[0] * ((j - g->lo[1]) + (long)g->n[1] * (k - g->lo[2]));
        if (back) p[I] = pI[J];
        else pI[J] = p[I];
      }
}

/* applypressure_kernel!  src/operators.jl:225-233 */
static void FN(applypressure_)(const FN(geo_) *g, REAL *u, const REAL *p) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = g->lo[2]; k < g->hi[2]; k++)
    for (int j = g->lo[1]; j < g->hi[1]; j++)
      for (int i = g->lo[0]; i < g->hi[0]; i++) {
        const long I = i + j * g->s[1] + k * g->s[2];
        for (int a = 0; a < g->D; a++) u[a * g->sc + I] -= (p[I + g->s[a]] - p[I]) / g->h[a];
      }
}

/* timestep!(::ExplicitRungeKuttaMethod) repeated nsteps times with fixed dt:
 * src/time_steppers/step_explicit_runge_kutta.jl:4-59, launch for launch (apply_bc_u!, fill!, convdiff,
 * force, closure, stage axpys incl. zero coefficients, apply_bc_u!, project!).  The Poisson solve of the
 * ghost-stripped array (FFTW in the reference) is delegated to `psolve` (pocketfft through scipy).
 * A, c: shifted tableau (s*s row-major, s).  theta = NaN: no closure.  force: stored field or NULL. */
int FN(cpuref_rk_steps_)(REAL *u, int D, const int *n, const double *h, double Re, double dt_, int nsteps, int s,
                         const double *A, const double *c, double theta_, const REAL *force, void (*psolve)(void *)) {
  FN(geo_) g;
  FN(geo_init_)(&g, D, n, h, Re);
  const long nsc = g.sc, nvec = g.sc * D, nint = (long)g.n[0] * g.n[1] * g.n[2];
  const int smag = theta_ == theta_;
  const REAL dt = (REAL)dt_, theta = (REAL)theta_;
  REAL *ustart = (REAL *)malloc(sizeof(REAL) * nvec), *p = (REAL *)calloc(nsc, sizeof(REAL));
  REAL *pI = (REAL *)malloc(sizeof(REAL) * nint), *sig = NULL, *sb = NULL;
  REAL **ku = (REAL **)malloc(sizeof(REAL *) * s);
  for (int i = 0; i < s; i++) ku[i] = (REAL *)malloc(sizeof(REAL) * nvec);
  if (smag) {
    sig = (REAL *)calloc(nsc * D * D, sizeof(REAL));
    sb = (REAL *)calloc(nvec, sizeof(REAL));
  }
  REAL vol = g.h[0] * g.h[1];
  if (D == 3) vol = vol * g.h[2];
  for (int step = 0; step < nsteps; step++) {
    FN(copy_)(ustart, u, nvec);                               /* :14 */
    for (int i = 0; i < s; i++) {
      FN(bc_fill_)(&g, u, D, 1);                              /* :19 */
      FN(fill_)(ku[i], nvec, (REAL)0);                        /* operators.jl:971 */
      FN(convdiff_)(&g, ku[i], u);                            /* :972 */
      if (force) FN(add_)(ku[i], force, nvec);                /* :973 */
      if (smag) {                                             /* :29, operators.jl:1300-1304 */
        FN(smagtensor_)(&g, sig, u, theta);
        FN(bc_fill_)(&g, sig, 1, D * D);
        FN(divoftensor_)(&g, sb, sig);
        FN(add_)(ku[i], sb, nvec);
      }
      FN(copy_)(u, ustart, nvec);                             /* :35 */
      for (int j = 0; j <= i; j++) FN(axpy_)(u, dt * (REAL)A[i * s + j], ku[j], nvec); /* :36-38 */
      FN(bc_fill_)(&g, u, D, 1);                              /* :48 */
      FN(divergence_)(&g, p, u);                              /* project! pressure.jl:73 */
      FN(scale_)(p, nsc, vol);                                /* :74 */
      FN(strip_)(&g, pI, p, 0);                               /* psolve!: pressure.jl:320 */
      psolve(pI);                                             /* :323-344 */
      FN(strip_)(&g, pI, p, 1);                               /* :347 */
      FN(bc_fill_)(&g, p, 1, 1);                              /* :78 */
      FN(applypressure_)(&g, u, p);                           /* :81 */
    }
    FN(bc_fill_)(&g, u, D, 1);                                /* :55 */
  }
  (void)c;
  free(ustart); free(p); free(pI);
  for (int i = 0; i < s; i++) free(ku[i]);
  free(ku);
  if (sig) free(sig);
  if (sb) free(sb);
  return 0;
}

#undef FN
#undef CAT
#undef CAT_
