// fast_stage_emul.cu -- runs the code of k_stage3d_fast (csrc/fast_stage.cuh) ON THE HOST: the FsThread objects of one
// block are stepped one after the other between the block barriers, TMA boxes are copied by hand with the kernel's own
// box coordinates, and the result (k_i, u', Poisson right-hand side) is compared with a plain periodic triple loop in
// double precision.  Checks the kernel's index arithmetic, slot rotation, carried registers, halo threads and chunk
// boundaries without a GPU (tests/test_fast_stage_emul.py builds and runs it; the GPU tests check the real launch).
//
//   nvcc -std=c++17 -O1 -gencode arch=compute_100a,code=sm_100a -o fast_stage_emul tools/fast_stage_emul.cu
//   ./fast_stage_emul [n0 n1 n2 zc]          exit code 0 = all variants agree
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../incompressiblenavierstokes.jl_b200/csrc/fast_stage.cuh"

using namespace insb200;

typedef float T;

struct Padded {  // the library's padded layout (ctx.cuh): 64 B of x margin, 2 / 3 ghost layers in y and z
  int n[3], org[3], ext[3];
  long long sy, sz, sc, corg;
  Padded(int n0, int n1, int n2) {
    n[0] = n0; n[1] = n1; n[2] = n2;
    org[0] = 64 / (int)sizeof(T); org[1] = 2; org[2] = 2;
    ext[0] = n0 + 128 / (int)sizeof(T); ext[1] = n1 + 5; ext[2] = n2 + 5;
    sy = ext[0]; sz = (long long)ext[0] * ext[1]; sc = sz * ext[2];
    corg = org[0] + org[1] * sy + org[2] * sz;
  }
  static int wrap(int i, int n) { return ((i % n) + n) % n; }
  std::vector<T> field(int nc) const { return std::vector<T>((size_t)sc * nc + 64, T(0)); }
  // interior value -> every periodic image inside the padded box
  void fill(std::vector<T> &f, int nc, const std::vector<double> &src) const {
    for (int c = 0; c < nc; c++)
      for (int z = 0; z < ext[2]; z++)
        for (int y = 0; y < ext[1]; y++)
          for (int x = 0; x < ext[0]; x++) {
            const int i = wrap(x - org[0], n[0]), j = wrap(y - org[1], n[1]), k = wrap(z - org[2], n[2]);
            f[(size_t)c * sc + z * sz + y * sy + x] = (T)src[((size_t)c * n[2] + k) * n[1] * n[0] + (size_t)j * n[0] + i];
          }
  }
};

struct Ref {  // periodic interior arrays in double, index (c, k, j, i)
  int n0, n1, n2;
  size_t at(int c, int i, int j, int k) const {
    return ((size_t)c * n2 + Padded::wrap(k, n2)) * n1 * n0 + (size_t)Padded::wrap(j, n1) * n0 + Padded::wrap(i, n0);
  }
};

template <typename O>
static bool same_offs(const O &a, const O &b) {
  return a.u_mid == b.u_mid && a.u_top == b.u_top && a.u_bm == b.u_bm && a.u_bt == b.u_bt && a.u_a == b.u_a && a.p_a == b.p_a &&
         a.p_b == b.p_b && a.g_w == b.g_w && a.g_m == b.g_m && a.g_t == b.g_t && a.stg == b.stg;
}

template <bool PGRAD, bool SMAG, bool GEN>
static int run_variant(int n0, int n1, int n2, int zc, bool force, int nkin, bool kout) {
  typedef FastCfg<T, SMAG> Cfg;
  Padded P(n0, n1, n2);
  Ref R{n0, n1, n2};
  const size_t nint = (size_t)n0 * n1 * n2;
  std::mt19937 rng(1234);
  std::uniform_real_distribution<double> U(-1.0, 1.0);
  auto rnd = [&](int nc, double amp) {
    std::vector<double> v(nint * nc);
    for (auto &x : v) x = (double)(T)(amp * U(rng));  // representable in T
    return v;
  };
  const double h[3] = {1.0 / n0, 0.75 / n1, 1.25 / n2};
  const double visc = 1.0 / 300.0, theta = 0.17, dt = 1e-3;
  std::vector<double> up = rnd(3, 1.0), pp = rnd(1, PGRAD ? 1e-3 : 0.0), u0 = rnd(3, 1.0), fo = rnd(3, 1.0);
  std::vector<std::vector<double>> kin(nkin);
  for (auto &k : kin) k = rnd(3, 1.0);
  double coef[8] = {0.3 * dt, -0.2 * dt, 0.7 * dt, 0.1 * dt}, coef_self = 0.5 * dt;

  // ---- reference: u = u' - grad p ; F ; sigma ; div sigma ; combine ; divergence
  std::vector<double> u(nint * 3), F(nint * 3), unew(nint * 3), dv(nint);
  const double rh[3] = {1 / h[0], 1 / h[1], 1 / h[2]};
  const int e[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int k = 0; k < n2; k++) for (int j = 0; j < n1; j++) for (int i = 0; i < n0; i++)
    for (int a = 0; a < 3; a++)
      u[R.at(a, i, j, k)] = up[R.at(a, i, j, k)] - (PGRAD ? (pp[R.at(0, i + e[a][0], j + e[a][1], k + e[a][2])] - pp[R.at(0, i, j, k)]) * rh[a] : 0.0);
  auto UU = [&](int c, int i, int j, int k) { return u[R.at(c, i, j, k)]; };
  std::vector<double> sig(nint * 9, 0.0);
  if (SMAG) {
    for (int k = 0; k < n2; k++) for (int j = 0; j < n1; j++) for (int i = 0; i < n0; i++) {
      double G[3][3];
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) {
        const int *ea = e[a], *eb = e[b];
        if (a == b) G[a][b] = (UU(a, i, j, k) - UU(a, i - ea[0], j - ea[1], k - ea[2])) * rh[a];
        else G[a][b] = ((UU(a, i + eb[0], j + eb[1], k + eb[2]) - UU(a, i, j, k)) * rh[b] +
                        (UU(a, i - ea[0] + eb[0], j - ea[1] + eb[1], k - ea[2] + eb[2]) - UU(a, i - ea[0], j - ea[1], k - ea[2])) * rh[b] +
                        (UU(a, i, j, k) - UU(a, i - eb[0], j - eb[1], k - eb[2])) * rh[b] +
                        (UU(a, i - ea[0], j - ea[1], k - ea[2]) - UU(a, i - ea[0] - eb[0], j - ea[1] - eb[1], k - ea[2] - eb[2])) * rh[b]) / 4;
      }
      double S[3][3], ss = 0;
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) { S[a][b] = 0.5 * (G[a][b] + G[b][a]); }
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) ss += S[a][b] * S[a][b];
      const double d2 = h[0] * h[0] + h[1] * h[1] + h[2] * h[2];
      const double nut = theta * theta * d2 * std::sqrt(2 * ss);
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) sig[R.at(a * 3 + b, i, j, k)] = 2 * nut * S[a][b];
    }
  }
  auto SG = [&](int a, int b, int i, int j, int k) { return sig[R.at(a * 3 + b, i, j, k)]; };
  for (int k = 0; k < n2; k++) for (int j = 0; j < n1; j++) for (int i = 0; i < n0; i++)
    for (int a = 0; a < 3; a++) {
      const int *ea = e[a];
      double f = 0;
      for (int b = 0; b < 3; b++) {
        const int *eb = e[b];
        const double uI = UU(a, i, j, k), um = UU(a, i - eb[0], j - eb[1], k - eb[2]), upl = UU(a, i + eb[0], j + eb[1], k + eb[2]);
        const double ua1 = (um + uI) / 2, ua2 = (uI + upl) / 2;
        const double ub1 = 0.5 * UU(b, i - eb[0], j - eb[1], k - eb[2]) + 0.5 * UU(b, i - eb[0] + ea[0], j - eb[1] + ea[1], k - eb[2] + ea[2]);
        const double ub2 = 0.5 * UU(b, i, j, k) + 0.5 * UU(b, i + ea[0], j + ea[1], k + ea[2]);
        f += (visc * ((upl - uI) * rh[b] - (uI - um) * rh[b]) - (ua2 * ub2 - ua1 * ub1)) * rh[b];
      }
      if (force) f += fo[R.at(a, i, j, k)];
      if (SMAG) {
        double s = 0;
        for (int b = 0; b < 3; b++) {
          const int *eb = e[b];
          double s1, s2;
          if (a == b) { s2 = SG(a, b, i + eb[0], j + eb[1], k + eb[2]); s1 = SG(a, b, i, j, k); }
          else {
            s2 = (SG(a, b, i, j, k) + SG(a, b, i + eb[0], j + eb[1], k + eb[2]) + SG(a, b, i + ea[0] + eb[0], j + ea[1] + eb[1], k + ea[2] + eb[2]) + SG(a, b, i + ea[0], j + ea[1], k + ea[2])) / 4;
            s1 = (SG(a, b, i - eb[0], j - eb[1], k - eb[2]) + SG(a, b, i, j, k) + SG(a, b, i + ea[0] - eb[0], j + ea[1] - eb[1], k + ea[2] - eb[2]) + SG(a, b, i + ea[0], j + ea[1], k + ea[2])) / 4;
          }
          s += (s2 - s1) * rh[b];
        }
        f += s;
      }
      F[R.at(a, i, j, k)] = f;
      double acc = PGRAD ? u0[R.at(a, i, j, k)] : UU(a, i, j, k);
      for (int q = 0; q < nkin; q++) acc += coef[q] * kin[q][R.at(a, i, j, k)];
      unew[R.at(a, i, j, k)] = acc + coef_self * f;
    }
  const double vol = h[0] * h[1] * h[2];
  for (int k = 0; k < n2; k++) for (int j = 0; j < n1; j++) for (int i = 0; i < n0; i++) {
    double d = 0;
    for (int a = 0; a < 3; a++) d += (unew[R.at(a, i, j, k)] - unew[R.at(a, i - e[a][0], j - e[a][1], k - e[a][2])]) * rh[a];
    dv[R.at(0, i, j, k)] = d * vol;
  }

  // ---- the kernel's code on the host
  std::vector<T> d_up = P.field(3), d_p = P.field(1), d_u0 = P.field(3), d_f = P.field(3), d_k = P.field(3), d_un = P.field(3), d_div = P.field(1);
  std::vector<std::vector<T>> d_kin(nkin);
  P.fill(d_up, 3, up); P.fill(d_p, 1, pp); P.fill(d_u0, 3, u0); P.fill(d_f, 3, fo);
  for (int q = 0; q < nkin; q++) { d_kin[q] = P.field(3); P.fill(d_kin[q], 3, kin[q]); }
  TmaStageArgs<T> A;
  memset(&A, 0, sizeof A);
  for (int c = 0; c < 3; c++) {
    const long long co = P.corg + c * P.sc;
    A.u0[c] = d_u0.data() + co;
    A.force[c] = force ? d_f.data() + co : nullptr;
    A.kout[c] = kout ? d_k.data() + co : nullptr;
    A.unew[c] = d_un.data() + co;
    for (int q = 0; q < nkin; q++) A.kin[q][c] = d_kin[q].data() + co;
  }
  A.div = d_div.data() + P.corg;
  A.nk = nkin;
  for (int q = 0; q < nkin; q++) A.coef[q] = (T)coef[q];
  A.coef_self = (T)coef_self;
  A.sy = P.sy; A.sz = P.sz; A.sc = P.sc;
  FastK<T> K;
  for (int a = 0; a < 3; a++) { K.rh[a] = (T)rh[a]; K.vrr[a] = (T)(visc * rh[a] * rh[a]); K.q[a] = (T)(0.25 * rh[a]); }
  K.vol = (T)vol;
  K.e2c = (T)(2 * theta * theta * (h[0] * h[0] + h[1] * h[1] + h[2] * h[2]));
  Geo g;
  g.n[0] = n0; g.n[1] = n1; g.n[2] = n2; g.wrap[0] = g.wrap[1] = g.wrap[2] = 1;

  typedef FsThread<T, PGRAD, SMAG, GEN> Thr;
  std::vector<T> smem(Cfg::ELEMS + 64);
  std::vector<Thr> th(Cfg::NT);
  for (int bz = 0; bz < (n2 + zc - 1) / zc; bz++)
    for (int by = 0; by < n1 / Cfg::TY; by++)
      for (int bx = 0; bx < n0 / Cfg::TX; bx++) {
        for (auto &v : smem) v = std::nanf("");  // anything read before it is written shows up
        for (int t = 0; t < Cfg::NT; t++) th[t].init(smem.data(), &A, &K, g, zc, t, t >> 5, bx, by, bz);
        const int kbeg = th[0].kbeg, nk = th[0].nk, qlast = th[0].qlast;
        const int x0 = P.org[0] + bx * Cfg::TX - Cfg::XL, y0 = P.org[1] + by * Cfg::TY - 2, z0 = P.org[2] + kbeg - 2;
        auto box = [&](T *dst, const std::vector<T> &src, int nc, int rows, int z) {  // cp.async.bulk.tensor.4d, box (BX, rows, 1, nc)
          for (int c = 0; c < nc; c++)
            for (int y = 0; y < rows; y++)
              for (int x = 0; x < Cfg::BX; x++) {
                if (x0 + x < 0 || x0 + x >= P.ext[0] || y0 + y < 0 || y0 + y >= P.ext[1] || z < 0 || z >= P.ext[2]) { printf("box out of bounds\n"); exit(2); }
                dst[(c * rows + y) * Cfg::BX + x] = src[(size_t)c * P.sc + (size_t)z * P.sz + (size_t)(y0 + y) * P.sy + x0 + x];
              }
        };
        auto issue = [&](int q) {
          if (PGRAD) {
            if (q == 0) box(smem.data() + Cfg::POFF, d_p, 1, Cfg::BYP, z0);
            box(smem.data() + Cfg::POFF + ((q + 1) % Cfg::NPS) * Cfg::PSLOT, d_p, 1, Cfg::BYP, z0 + q + 1);
          }
          box(smem.data() + (q % Cfg::NS) * Cfg::USLOT, d_up, 3, Cfg::BY, z0 + q);
        };
        for (int q = 0; q < Cfg::NS - 1; q++) if (q <= qlast) issue(q);
        int r = Cfg::R0;
        for (; r < 0; r++) {
          const FsOffs<Cfg> S = FsOffs<Cfg>::make(r);
          for (int t = 0; t < Cfg::NT; t++) { if (th[t].cell_thr) th[t].template pre<false, true>(r, S); else th[t].template pre<false, false>(r, S); }
          if (r == Cfg::R0 && Cfg::NS - 1 <= qlast) issue(Cfg::NS - 1);
          if (r >= -2 && r + 2 + Cfg::NS <= qlast) issue(r + 2 + Cfg::NS);
          for (int t = 0; t < Cfg::NT; t++) { if (th[t].cell_thr) th[t].template post<false, true>(r, S); else th[t].template post<false, false>(r, S); }
        }
        int H = 0, HP = 0;
        auto one = [&](auto jc) {
          constexpr int J = decltype(jc)::value;
          const FsOffs<Cfg> S = FsOffs<Cfg>::template steady<J>(H, HP);
          const FsOffs<Cfg> G = FsOffs<Cfg>::make(r);  // the unrolled offsets must equal the general ones
          if (!same_offs(S, G)) { printf("steady offsets differ at r = %d\n", r); exit(3); }
          for (int t = 0; t < Cfg::NT; t++) { if (th[t].cell_thr) th[t].template pre<true, true>(r, S); else th[t].template pre<true, false>(r, S); }
          if (r + 2 + Cfg::NS <= qlast) issue(r + 2 + Cfg::NS);
          for (int t = 0; t < Cfg::NT; t++) { if (th[t].cell_thr) th[t].template post<true, true>(r, S); else th[t].template post<true, false>(r, S); }
          r++;
        };
        for (; r + 4 <= nk;) {
          one(IntC<0>{}); one(IntC<1>{}); one(IntC<2>{}); one(IntC<3>{});
          H = 4 * Cfg::USLOT - H;
          HP = 4 * Cfg::PSLOT - HP;
        }
        for (; r < nk; r++) {
          const FsOffs<Cfg> S = FsOffs<Cfg>::make(r);
          for (int t = 0; t < Cfg::NT; t++) { if (th[t].cell_thr) th[t].template pre<true, true>(r, S); else th[t].template pre<true, false>(r, S); }
          if (r + 2 + Cfg::NS <= qlast) issue(r + 2 + Cfg::NS);
          for (int t = 0; t < Cfg::NT; t++) { if (th[t].cell_thr) th[t].template post<true, true>(r, S); else th[t].template post<true, false>(r, S); }
        }
      }

  // ---- compare
  auto cmp = [&](const char *name, const std::vector<T> &got, int nc, const std::vector<double> &ref) {
    double emax = 0, rmax = 0;
    int bad[4] = {-1, -1, -1, -1};
    for (int c = 0; c < nc; c++) for (int k = 0; k < n2; k++) for (int j = 0; j < n1; j++) for (int i = 0; i < n0; i++) {
      const double a = got[(size_t)c * P.sc + P.corg + k * P.sz + j * P.sy + i], b = ref[R.at(c, i, j, k)];
      const double er = std::isfinite(a) ? std::fabs(a - b) : 1e30;
      if (er > emax) { emax = er; bad[0] = c; bad[1] = i; bad[2] = j; bad[3] = k; }
      rmax = std::max(rmax, std::fabs(b));
    }
    const bool ok = emax <= 3e-5 * rmax;
    printf("    %-5s max|err| = %.3e (max|ref| = %.3e) at c=%d i=%d j=%d k=%d  %s\n", name, emax, rmax, bad[0], bad[1], bad[2], bad[3], ok ? "ok" : "FAIL");
    return ok ? 0 : 1;
  };
  int fails = 0;
  printf("  PGRAD=%d SMAG=%d GEN=%d force=%d nkin=%d kout=%d\n", PGRAD, SMAG, GEN, force, nkin, kout);
  if (kout) fails += cmp("k_i", d_k, 3, F);
  fails += cmp("u'", d_un, 3, unew);
  fails += cmp("div", d_div, 1, dv);
  return fails;
}

int main(int argc, char **argv) {
  int n0 = 64, n1 = 32, n2 = 12, zc = 5;
  if (argc >= 5) { n0 = atoi(argv[1]); n1 = atoi(argv[2]); n2 = atoi(argv[3]); zc = atoi(argv[4]); }
  printf("fast_stage_emul: grid %d x %d x %d, z-chunks of %d planes\n", n0, n1, n2, zc);
  int fails = 0;
  fails += run_variant<false, false, false>(n0, n1, n2, zc, false, 0, true);   // first stage
  fails += run_variant<true, false, false>(n0, n1, n2, zc, false, 0, true);    // middle stage
  fails += run_variant<true, false, true>(n0, n1, n2, zc, true, 0, true);      // middle stage + force (GEN kernel)
  fails += run_variant<true, false, true>(n0, n1, n2, zc, false, 3, false);    // last RK44 row
  fails += run_variant<false, true, false>(n0, n1, n2, zc, false, 0, true);    // closure, first stage
  fails += run_variant<true, true, false>(n0, n1, n2, zc, false, 0, true);     // closure, middle stage
  fails += run_variant<true, true, true>(n0, n1, n2, zc, true, 0, true);       // closure, middle stage + force (GEN kernel)
  fails += run_variant<true, true, true>(n0, n1, n2, zc, false, 2, true);      // closure, general row with 2 earlier k_j
  printf(fails ? "FAILED (%d)\n" : "all variants agree\n", fails);
  return fails != 0;
}
