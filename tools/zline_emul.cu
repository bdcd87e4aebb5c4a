// zline_emul.cu -- runs the data routines of csrc/zline.cuh (the transpose-free last-axis solve of the slab-decomposed
// pressure solver) on the CPU for P emulated ranks with C sub-chunks each, pass by pass as the kernels would, and compares
// with the reference's definition: transform along z, divide by -(ax+ay+az), zero the constant mode, transform back
// (src/pressure.jl:323-344), evaluated mode by mode as a plain O(n^2) DFT in double precision.
//   usage: zline_emul nx ny n P C   (ny = 1: the 2-D case, one row per "plane")
// Built by tests/test_zline_emul.py with nvcc (host execution only; no GPU involved).
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../incompressiblenavierstokes.jl_b200/csrc/zline.cuh"

using namespace insb200;

static double sinpi01(double x) { return std::sin(M_PI * x); }

template <typename T>
static double run(int nx, int ny, int n, int P, int C, bool verbose) {
  const int D = ny > 1 ? 3 : 2;
  const int nxh = nx / 2 + 1, pitch = (nxh + 7) & ~7, nky = ny;
  const double L[3] = {1.0, 2.0, 1.5};
  // grid spacings and the symbol tables as upload_symbol (insb200.cu) forms them
  double h[3] = {L[0] / nx, D == 3 ? L[1] / ny : L[2] / n, L[2] / n};
  T om = T(h[0]);
  for (int a = 1; a < D; a++) om = om * T(h[a]);
  auto table = [&](int a, int cnt, int npts) {
    std::vector<T> v(cnt);
    const T dx = T(h[a]);
    const double num = (double)(T(4) * om), den = (double)(dx * dx);
    for (int k = 0; k < cnt; k++) { const double s = sinpi01((double)k / npts); v[k] = (T)(num * s * s / den); }
    return v;
  };
  std::vector<T> ax = table(0, nxh, nx), ay, az = table(D - 1, n, n);
  if (D == 3) ay = table(1, ny, ny);
  const T dl = T(h[D - 1]);
  ZlGeo g;
  g.nkx = nxh; g.pitch = pitch; g.nky = nky; g.pstride = (long long)pitch * nky;
  g.mloc = n / P; g.C = C; g.mv = g.mloc / C; g.P = P; g.n = n;
  g.kappa = (double)(dl * dl) / (double)om;
  const double nin = (double)nx * (D == 3 ? ny : 1);
  g.scale = g.kappa / nin;
  g.xoff = 0; g.rank = 0;
  const long long area = zl_area_words(g);
  g.xoff = area;  // the second of the two areas, to exercise the offset
  std::mt19937_64 rng(7);
  std::normal_distribution<double> nd;
  std::vector<std::vector<Cx<T>>> S(P), X(P);
  std::vector<std::vector<std::complex<double>>> H(P);
  for (int r = 0; r < P; r++) {
    S[r].assign((size_t)g.pstride * g.mloc, Cx<T>{T(777), T(777)});
    X[r].assign((size_t)2 * area, Cx<T>{T(-1e30), T(1e30)});
    for (int j = 0; j < g.mloc; j++)
      for (int ky = 0; ky < nky; ky++)
        for (int kx = 0; kx < nxh; kx++) S[r][(size_t)j * g.pstride + (size_t)ky * pitch + kx] = Cx<T>{(T)nd(rng), (T)nd(rng)};
    H[r].resize(S[r].size());
    for (size_t i = 0; i < S[r].size(); i++) H[r][i] = {(double)S[r][i].x, (double)S[r][i].y};
  }
  PeerPtrs px;
  for (int r = 0; r < 16; r++) px.p[r] = r < P ? (void *)X[r].data() : nullptr;
  const T *ayp = D == 3 ? ay.data() : nullptr;
  for (int r = 0; r < P; r++) { g.rank = r; for (int c = 0; c < C; c++) for (long long i = 0; i < g.pstride; i++) zl_passA<T>(g, px, S[r].data(), ax.data(), ayp, i, c); }
  // (barrier)
  for (int r = 0; r < P; r++) {  // the constant in-plane mode, as k_zline_zero does it
    g.rank = r;
    std::vector<double2> col(n);
    double mx = 0, my = 0;
    for (int i = 0; i < n; i++) {
      const int q = i / g.mloc, jl = i - q * g.mloc;
      const Cx<T> v = X[q][g.xoff + (long long)C * 2 * g.pstride + jl];
      col[i] = double2{(double)v.x, (double)v.y};
      mx += col[i].x; my += col[i].y;
    }
    mx /= n; my /= n;
    for (int jl = 0; jl < g.mloc; jl++) {
      double sx = 0, sy = 0;
      for (int lane = 0; lane < 32; lane++) {  // the lanes of the warp that owns this point
        double ax_, ay_;
        zl_zero_point<double2>(col.data(), n, r * g.mloc + jl, lane, 32, mx, my, ax_, ay_);
        sx += ax_; sy += ay_;
      }
      S[r][(size_t)jl * g.pstride] = Cx<T>{(T)(g.scale * sx), (T)(g.scale * sy)};
    }
  }
  for (int r = 0; r < P; r++) { g.rank = r; for (int c = 0; c < C; c++) for (long long i = 0; i < g.pstride; i++) zl_passB<T>(g, px, S[r].data(), ax.data(), ayp, i, c); }
  // reference: per mode, DFT along the line in double
  double err2 = 0, ref2 = 0, emax = 0, rmax = 0;
  std::vector<std::complex<double>> f(n), F(n), p(n);
  const double invN = 1.0 / (nin * n);
  for (int ky = 0; ky < nky; ky++)
    for (int kx = 0; kx < nxh; kx++) {
      const size_t mo = (size_t)ky * pitch + kx;
      for (int j = 0; j < n; j++) f[j] = H[j / g.mloc][(size_t)(j % g.mloc) * g.pstride + mo];
      const double axy = D == 3 ? (double)(T)(ax[kx] + ay[ky]) : (double)ax[kx];
      for (int k = 0; k < n; k++) {
        std::complex<double> s = 0;
        for (int j = 0; j < n; j++) s += f[j] * std::polar(1.0, -2.0 * M_PI * (double)((long long)j * k % n) / n);
        const double sk = sinpi01((double)k / n);
        const double azk = (double)(T(4) * om) * sk * sk / (double)(dl * dl);
        F[k] = (kx == 0 && ky == 0 && k == 0) ? 0.0 : -s / (axy + azk) * invN;
      }
      for (int j = 0; j < n; j++) {
        std::complex<double> s = 0;
        for (int k = 0; k < n; k++) s += F[k] * std::polar(1.0, 2.0 * M_PI * (double)((long long)j * k % n) / n);
        const Cx<T> got = S[j / g.mloc][(size_t)(j % g.mloc) * g.pstride + mo];
        const double dx_ = (double)got.x - s.real(), dy_ = (double)got.y - s.imag();
        err2 += dx_ * dx_ + dy_ * dy_; ref2 += std::norm(s);
        emax = std::max(emax, std::hypot(dx_, dy_)); rmax = std::max(rmax, std::abs(s));
      }
    }
  // pitch padding must be untouched
  for (int r = 0; r < P; r++)
    for (int j = 0; j < g.mloc; j++)
      for (int ky = 0; ky < nky; ky++)
        for (int kx = nxh; kx < pitch; kx++)
          if (S[r][(size_t)j * g.pstride + (size_t)ky * pitch + kx].x != T(777)) { printf("padding overwritten\n"); return 1.0; }
  const double rel = std::sqrt(err2 / ref2);
  if (verbose) printf("  %s nx=%d ny=%d n=%d P=%d C=%d: rel l2 %.3e  max %.3e (of %.3e)\n", sizeof(T) == 4 ? "f32" : "f64", nx, ny, n, P, C, rel, emax, rmax);
  return std::max(rel, emax / rmax);
}

int main(int argc, char **argv) {
  if (argc < 6) { printf("usage: zline_emul nx ny n P C\n"); return 2; }
  const int nx = atoi(argv[1]), ny = atoi(argv[2]), n = atoi(argv[3]), P = atoi(argv[4]), C = atoi(argv[5]);
  if (n % (P * C)) { printf("n must be a multiple of P*C\n"); return 2; }
  const double e64 = run<double>(nx, ny, n, P, C, true);
  const double e32 = run<float>(nx, ny, n, P, C, true);
  const bool ok = e64 < 1e-11 && e32 < 5e-6;
  printf(ok ? "line solve agrees\n" : "MISMATCH\n");
  return ok ? 0 : 1;
}
