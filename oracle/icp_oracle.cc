// icp_oracle.cc — CPU restatement of the reference's ICP hot path (src/cpu.cc, src/load.cc).
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product path (icp_b200/, include/, the `icp` driver)
// may call, link or import this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, and only as the checker / the timed CPU baseline.
//
// Why a restatement: the reference's src/cpu.cc needs Eigen (fetched from GitLab at configure time,
// commit bcbaad6d874d451817457ae0603f953cda3c0c06, lib/CMakeLists.txt:3-6), which is neither vendored
// in /root/reference nor installed here, and there is no network.  Each function below cites the
// reference file:line it follows; Eigen facilities are restated from their documented behaviour.
//
// Parity pinning: the reference holds exactly one known answer for this path — `./icp cow_ref.txt
// cow_tr1.txt 10` runs 7 iterations (benchmark/callgrind.out.76685:5,28306-28308).  tests/ assert
// it.  Beyond that datum this restatement is checked against the reference's OWN src/cpu.cc, compiled in
// place against an Eigen stand-in (oracle/_ref, oracle/eigen_shim, tests/test_ref_cpu.py): indices, Horn
// solve, residuals and the 7-iteration loop agree.  Still "parity unpinned": Eigen's own numerics
// (EigenSolver's eigenvalue order, reduction order of long sums), which no stand-in can reproduce.
//
// Modes
//   nn_mode 0  literal : sqrt(pow(dx,2)+pow(dy,2)+pow(dz,2)), libm pow/sqrt   (src/cpu.cc:17-19)
//   nn_mode 1  mulsqrt : sqrt((dx*dx+dy*dy)+dz*dz)    — what the CUDA exact re-rank evaluates
//   nn_mode 2  mul     : (dx*dx+dy*dy)+dz*dz          — fastest CPU form
//   eig_mode 0 intended: eigenvector of the LARGEST eigenvalue of Horn's N (symmetric Jacobi)
//   eig_mode 1 literal : general real eigen-solver (Householder-Hessenberg + Francis double-shift
//                        QR, the published EISPACK hqr / JAMA algorithm that Eigen::EigenSolver
//                        derives from) + the reference's max_element_index rule incl. its bug
//                        (src/cpu.cc:81-91: `max` never updated).
//
// Build: see oracle/Makefile  (g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ------------------------------------------------------------------------------------------------
// 4x4 helpers
// ------------------------------------------------------------------------------------------------

// Symmetric cyclic Jacobi; returns eigenvalues w[4] and eigenvectors as columns of V (row-major).
void jacobi4(const double Ain[16], double w[4], double V[16])
{
    double A[16];
    std::memcpy(A, Ain, sizeof(A));
    for (int i = 0; i < 16; ++i) V[i] = (i % 5 == 0) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 64; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < 4; ++i)
            for (int j = i + 1; j < 4; ++j) off += A[i * 4 + j] * A[i * 4 + j];
        if (off == 0.0) break;
        for (int p = 0; p < 3; ++p)
            for (int q = p + 1; q < 4; ++q) {
                double apq = A[p * 4 + q];
                if (apq == 0.0) continue;
                double app = A[p * 4 + p], aqq = A[q * 4 + q];
                double theta = (aqq - app) / (2.0 * apq);
                double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; ++k) {  // A <- A * J
                    double akp = A[k * 4 + p], akq = A[k * 4 + q];
                    A[k * 4 + p] = c * akp - s * akq;
                    A[k * 4 + q] = s * akp + c * akq;
                }
                for (int k = 0; k < 4; ++k) {  // A <- J^T * A
                    double apk = A[p * 4 + k], aqk = A[q * 4 + k];
                    A[p * 4 + k] = c * apk - s * aqk;
                    A[q * 4 + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 4; ++k) {
                    double vkp = V[k * 4 + p], vkq = V[k * 4 + q];
                    V[k * 4 + p] = c * vkp - s * vkq;
                    V[k * 4 + q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < 4; ++i) w[i] = A[i * 4 + i];
}

// General real eigen-solver for an n x n matrix (n = 4 here), real spectrum assumed (N is symmetric).
// Householder reduction to Hessenberg form followed by the Francis double-shift QR iteration on the
// Hessenberg matrix with accumulation of the transformations, then back-substitution for the
// eigenvectors of the quasi-triangular factor — the textbook hqr2 procedure.  Eigenvalues are read
// off the diagonal top to bottom, which is the order Eigen::EigenSolver::eigenvalues() reports and
// hence what the reference's max_element_index sees (src/cpu.cc:128-131).
struct GeneralEig4 {
    static constexpr int n = 4;
    double H[4][4];   // Hessenberg -> quasi-triangular T
    double Z[4][4];   // accumulated orthogonal transforms
    double wr[4];
    double vec[4][4]; // eigenvectors as columns, unit 2-norm

    void hessenberg(const double A[16])
    {
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) { H[i][j] = A[i * 4 + j]; Z[i][j] = (i == j); }
        for (int k = 0; k < n - 2; ++k) {
            // Householder vector annihilating H[k+2..n-1][k]
            double alpha = 0.0;
            for (int i = k + 1; i < n; ++i) alpha += H[i][k] * H[i][k];
            double tail = alpha - H[k + 1][k] * H[k + 1][k];
            if (tail <= 0.0) continue;
            alpha = std::sqrt(alpha);
            double beta = (H[k + 1][k] >= 0.0) ? -alpha : alpha;
            double v[4] = {0, 0, 0, 0};
            v[k + 1] = H[k + 1][k] - beta;
            for (int i = k + 2; i < n; ++i) v[i] = H[i][k];
            double vnorm2 = 0.0;
            for (int i = k + 1; i < n; ++i) vnorm2 += v[i] * v[i];
            if (vnorm2 == 0.0) continue;
            double tau = 2.0 / vnorm2;
            // H <- (I - tau v v^T) H
            for (int j = 0; j < n; ++j) {
                double s = 0.0;
                for (int i = k + 1; i < n; ++i) s += v[i] * H[i][j];
                s *= tau;
                for (int i = k + 1; i < n; ++i) H[i][j] -= s * v[i];
            }
            // H <- H (I - tau v v^T);  Z <- Z (I - tau v v^T)
            for (int i = 0; i < n; ++i) {
                double s = 0.0, sz = 0.0;
                for (int j = k + 1; j < n; ++j) { s += H[i][j] * v[j]; sz += Z[i][j] * v[j]; }
                s *= tau; sz *= tau;
                for (int j = k + 1; j < n; ++j) { H[i][j] -= s * v[j]; Z[i][j] -= sz * v[j]; }
            }
            for (int i = k + 2; i < n; ++i) H[i][k] = 0.0;
        }
    }

    // One Householder reflector of size `sz` (2 or 3) acting on rows/cols k..k+sz-1.
    void reflect(int k, int sz, const double v_in[3], int il, int iu)
    {
        (void)il; (void)iu;
        double x = v_in[0];
        double tail2 = 0.0;
        for (int i = 1; i < sz; ++i) tail2 += v_in[i] * v_in[i];
        if (tail2 == 0.0) return;
        double beta = std::sqrt(x * x + tail2);
        if (x >= 0.0) beta = -beta;
        double ess[3] = {1.0, 0.0, 0.0};
        for (int i = 1; i < sz; ++i) ess[i] = v_in[i] / (x - beta);
        double tau = (beta - x) / beta;
        for (int j = 0; j < n; ++j) {  // left: rows k..k+sz-1
            double s = 0.0;
            for (int i = 0; i < sz; ++i) s += ess[i] * H[k + i][j];
            s *= tau;
            for (int i = 0; i < sz; ++i) H[k + i][j] -= s * ess[i];
        }
        for (int i = 0; i < n; ++i) {  // right: cols k..k+sz-1
            double s = 0.0, sz_ = 0.0;
            for (int j = 0; j < sz; ++j) { s += H[i][k + j] * ess[j]; sz_ += Z[i][k + j] * ess[j]; }
            s *= tau; sz_ *= tau;
            for (int j = 0; j < sz; ++j) { H[i][k + j] -= s * ess[j]; Z[i][k + j] -= sz_ * ess[j]; }
        }
    }

    bool schur()
    {
        const double eps = 2.220446049250313e-16;
        double norm = 0.0;
        for (int j = 0; j < n; ++j)
            for (int i = 0; i <= std::min(n - 1, j + 1); ++i) norm += std::fabs(H[i][j]);
        if (norm == 0.0) return true;
        double tiny = std::max(norm * eps * eps, 2.2250738585072014e-308);
        int iu = n - 1, iter = 0, total = 0;
        double exshift = 0.0;
        while (iu >= 0) {
            int il = iu;
            while (il > 0) {
                double s = std::fabs(H[il - 1][il - 1]) + std::fabs(H[il][il]);
                s = std::max(s * eps, tiny);
                if (std::fabs(H[il][il - 1]) <= s) break;
                --il;
            }
            if (il == iu) {                       // one real root
                H[iu][iu] += exshift;
                if (iu > 0) H[iu][iu - 1] = 0.0;
                --iu; iter = 0;
            } else if (il == iu - 1) {            // 2x2 block
                double p = 0.5 * (H[iu - 1][iu - 1] - H[iu][iu]);
                double q = p * p + H[iu][iu - 1] * H[iu - 1][iu];
                H[iu][iu] += exshift; H[iu - 1][iu - 1] += exshift;
                if (q >= 0.0) {                   // two real roots: rotate the block to triangular form
                    double z = std::sqrt(std::fabs(q));
                    double a = (p >= 0.0) ? p + z : p - z, b = H[iu][iu - 1];
                    double r = std::hypot(a, b);
                    double c = a / r, s = b / r;  // Givens G = [c s; -s c] with G [a;b] = [r;0]
                    for (int j = iu - 1; j < n; ++j) {   // left: rows iu-1, iu
                        double h1 = H[iu - 1][j], h2 = H[iu][j];
                        H[iu - 1][j] = c * h1 + s * h2;
                        H[iu][j] = -s * h1 + c * h2;
                    }
                    for (int i = 0; i <= iu; ++i) {      // right: cols iu-1, iu
                        double h1 = H[i][iu - 1], h2 = H[i][iu];
                        H[i][iu - 1] = c * h1 + s * h2;
                        H[i][iu] = -s * h1 + c * h2;
                    }
                    for (int i = 0; i < n; ++i) {
                        double z1 = Z[i][iu - 1], z2 = Z[i][iu];
                        Z[i][iu - 1] = c * z1 + s * z2;
                        Z[i][iu] = -s * z1 + c * z2;
                    }
                    H[iu][iu - 1] = 0.0;
                } else {
                    return false;                 // complex pair: cannot happen for symmetric N
                }
                if (iu > 1) H[iu - 1][iu - 2] = 0.0;
                iu -= 2; iter = 0;
            } else {                              // Francis double-shift step on rows il..iu
                double sh0 = H[iu][iu], sh1 = H[iu - 1][iu - 1], sh2 = H[iu][iu - 1] * H[iu - 1][iu];
                if (iter == 10) {                 // Wilkinson's exceptional shift
                    exshift += sh0;
                    for (int i = 0; i <= iu; ++i) H[i][i] -= sh0;
                    double s = std::fabs(H[iu][iu - 1]) + std::fabs(H[iu - 1][iu - 2]);
                    sh0 = sh1 = 0.75 * s; sh2 = -0.4375 * s * s;
                }
                if (iter == 30) {                 // MATLAB's exceptional shift
                    double s = (sh1 - sh0) / 2.0;
                    s = s * s + sh2;
                    if (s > 0.0) {
                        s = std::sqrt(s);
                        if (sh1 < sh0) s = -s;
                        s = s + (sh1 - sh0) / 2.0;
                        s = sh0 - sh2 / s;
                        exshift += s;
                        for (int i = 0; i <= iu; ++i) H[i][i] -= s;
                        sh0 = sh1 = sh2 = 0.964;
                    }
                }
                ++iter; ++total;
                if (total > 40 * n) return false;
                int im; double v[3];
                for (im = iu - 2; im >= il; --im) {
                    double Tmm = H[im][im], r = sh0 - Tmm, s = sh1 - Tmm;
                    v[0] = (r * s - sh2) / H[im + 1][im] + H[im][im + 1];
                    v[1] = H[im + 1][im + 1] - Tmm - r - s;
                    v[2] = H[im + 2][im + 1];
                    if (im == il) break;
                    double lhs = H[im][im - 1] * (std::fabs(v[1]) + std::fabs(v[2]));
                    double rhs = v[0] * (std::fabs(H[im - 1][im - 1]) + std::fabs(Tmm) + std::fabs(H[im + 1][im + 1]));
                    if (std::fabs(lhs) < eps * rhs) break;
                }
                for (int k = im; k <= iu - 2; ++k) {
                    bool first = (k == im);
                    double w[3];
                    if (first) { w[0] = v[0]; w[1] = v[1]; w[2] = v[2]; }
                    else { w[0] = H[k][k - 1]; w[1] = H[k + 1][k - 1]; w[2] = H[k + 2][k - 1]; }
                    double tail2 = w[1] * w[1] + w[2] * w[2];
                    if (tail2 == 0.0) continue;
                    double beta = std::sqrt(w[0] * w[0] + tail2);
                    if (w[0] >= 0.0) beta = -beta;
                    if (first && k != il) H[k][k - 1] = -H[k][k - 1];
                    else if (!first) H[k][k - 1] = beta;
                    double ess1 = w[1] / (w[0] - beta), ess2 = w[2] / (w[0] - beta);
                    double tau = (beta - w[0]) / beta;
                    for (int j = k; j < n; ++j) {
                        double s = H[k][j] + ess1 * H[k + 1][j] + ess2 * H[k + 2][j];
                        s *= tau;
                        H[k][j] -= s; H[k + 1][j] -= s * ess1; H[k + 2][j] -= s * ess2;
                    }
                    int rmax = std::min(iu, k + 3);
                    for (int i = 0; i <= rmax; ++i) {
                        double s = H[i][k] + ess1 * H[i][k + 1] + ess2 * H[i][k + 2];
                        s *= tau;
                        H[i][k] -= s; H[i][k + 1] -= s * ess1; H[i][k + 2] -= s * ess2;
                    }
                    for (int i = 0; i < n; ++i) {
                        double s = Z[i][k] + ess1 * Z[i][k + 1] + ess2 * Z[i][k + 2];
                        s *= tau;
                        Z[i][k] -= s; Z[i][k + 1] -= s * ess1; Z[i][k + 2] -= s * ess2;
                    }
                }
                {   // final 2-row reflector on rows iu-1, iu
                    int k = iu - 1;
                    double w0 = H[k][k - 1], w1 = H[k + 1][k - 1];
                    if (w1 != 0.0) {
                        double beta = std::sqrt(w0 * w0 + w1 * w1);
                        if (w0 >= 0.0) beta = -beta;
                        double ess1 = w1 / (w0 - beta), tau = (beta - w0) / beta;
                        H[k][k - 1] = beta;
                        for (int j = k; j < n; ++j) {
                            double s = (H[k][j] + ess1 * H[k + 1][j]) * tau;
                            H[k][j] -= s; H[k + 1][j] -= s * ess1;
                        }
                        for (int i = 0; i <= iu; ++i) {
                            double s = (H[i][k] + ess1 * H[i][k + 1]) * tau;
                            H[i][k] -= s; H[i][k + 1] -= s * ess1;
                        }
                        for (int i = 0; i < n; ++i) {
                            double s = (Z[i][k] + ess1 * Z[i][k + 1]) * tau;
                            Z[i][k] -= s; Z[i][k + 1] -= s * ess1;
                        }
                    }
                }
                for (int i = im + 2; i <= iu; ++i) {   // clean up below the sub-diagonal
                    H[i][i - 2] = 0.0;
                    if (i > im + 2) H[i][i - 3] = 0.0;
                }
            }
        }
        return true;
    }

    // Back-substitute eigenvectors of upper-triangular T, map through Z, normalise columns.
    void vectors()
    {
        const double eps = 2.220446049250313e-16;
        double norm = 0.0;
        for (int j = 0; j < n; ++j)
            for (int i = 0; i <= j; ++i) norm += std::fabs(H[i][j]);
        double X[4][4] = {{0}};
        for (int k = n - 1; k >= 0; --k) {
            wr[k] = H[k][k];
            X[k][k] = 1.0;
            for (int i = k - 1; i >= 0; --i) {
                double r = 0.0;
                for (int j = i + 1; j <= k; ++j) r += H[i][j] * X[j][k];
                double w = H[i][i] - wr[k];
                if (w == 0.0) w = eps * norm;
                X[i][k] = -r / w;
            }
        }
        for (int k = 0; k < n; ++k) {
            double nn = 0.0;
            for (int i = 0; i < n; ++i) {
                double s = 0.0;
                for (int j = 0; j <= k; ++j) s += Z[i][j] * X[j][k];
                vec[i][k] = s; nn += s * s;
            }
            nn = std::sqrt(nn);
            for (int i = 0; i < n; ++i) vec[i][k] /= nn;
        }
    }

    bool compute(const double A[16])
    {
        hessenberg(A);
        if (!schur()) return false;
        vectors();
        return true;
    }
};

// src/cpu.cc:81-91 — note `max` is never updated: returns the LAST i in {1,2,3} with
// eigenvalue(i) > eigenvalue(0), or 0.
int max_element_index_literal(const double ev[4])
{
    int index = 0;
    double max = ev[0];
    for (int i = 1; i < 4; i++)
        if (ev[i] > max) index = i;
    return index;
}

// libm's pow, reached through a volatile pointer: g++ folds pow(x, 2.0) with a visible constant exponent into x * x
// (no call is left at -O1 and above), whereas the reference's Eigen expression array().pow(2) (src/cpu.cc:17) does
// call libm's pow with a run-time exponent - 176 975 589 calls in benchmark/callgrind.out.76685 - and glibc's
// pow(x, 2) differs from x * x by one ulp for ~0.08 % of inputs.  The literal mode must see that.
double (*volatile g_libm_pow)(double, double) = ::pow;

inline double dist_eval(int mode, double px, double py, double pz, double mx, double my, double mz)
{
    double dx = px - mx, dy = py - my, dz = pz - mz;      // src/cpu.cc:16 (pi - mk)
    if (mode == 0) {                                      // :17-18
        double (*const lp)(double, double) = g_libm_pow;
        return std::sqrt((lp(dx, 2.0) + lp(dy, 2.0)) + lp(dz, 2.0));
    }
    double s = (dx * dx + dy * dy) + dz * dz;
    return mode == 1 ? std::sqrt(s) : s;
}

}  // namespace

extern "C" {

int oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// src/load.cc:3-33.  n = (#lines) - 1; header skipped; sscanf("%lf,%lf,%lf") takes the first three
// comma-separated numbers of each line (extra columns ignored, missing ones stay 0).
// Returns n, or -2 if the file cannot be opened (the reference prints and exit(2)s, load.cc:11-14).
// xyz_aos may be NULL to query n; otherwise it receives min(n, cap) points as x0 y0 z0 x1 ... — the
// in-memory layout of the reference's column-major 3 x n MatrixXd.
long oracle_load_matrix(const char* filename, double* xyz_aos, long cap)
{
    std::ifstream infile(filename);
    if (!infile.is_open()) return -2;
    std::string line;
    long n = 0;
    while (std::getline(infile, line)) n++;
    n--;
    if (!xyz_aos) return n;
    infile.clear();
    infile.seekg(0, std::ios::beg);
    std::getline(infile, line);
    for (long i = 0; i < n && i < cap; i++) {
        std::getline(infile, line);
        double x = 0., y = 0., z = 0.;
        std::sscanf(line.c_str(), "%lf,%lf,%lf", &x, &y, &z);
        xyz_aos[3 * i] = x; xyz_aos[3 * i + 1] = y; xyz_aos[3 * i + 2] = z;
    }
    return n;
}

// src/load.cc:83-97: header "Points_0,Points_1,Points_2", then x,y,z per line with the default
// ostream formatting (6 significant digits, %g-style).
int oracle_write_matrix(const char* filename, const double* xyz_aos, long n)
{
    std::ofstream output(filename);
    if (!output.is_open()) return -1;
    output << "Points_0,Points_1,Points_2" << std::endl;
    for (long i = 0; i < n; i++)
        output << xyz_aos[3 * i] << ',' << xyz_aos[3 * i + 1] << ',' << xyz_aos[3 * i + 2] << std::endl;
    return 0;
}

// src/cpu.cc:5-27.  For each scene column j the first (lowest-index) minimum over k of the
// distance expression; Y[:,j] = m[:,k*].  idx (int32, nullable) exposes k*, which the reference
// never returns.  OpenMP over j only: per-point results are unaffected.
void oracle_closest_matrix(long np, long nm, const double* p, const double* m, int nn_mode,
                           int nthreads, int32_t* idx, double* Y)
{
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 64) num_threads(nthreads)
#endif
    for (long j = 0; j < np; j++) {
        const double px = p[3 * j], py = p[3 * j + 1], pz = p[3 * j + 2];
        double best = 0.0; long bk = -1;
        if (nn_mode == 0) {
            for (long k = 0; k < nm; k++) {
                double d = dist_eval(0, px, py, pz, m[3 * k], m[3 * k + 1], m[3 * k + 2]);
                if (bk < 0 || d < best) { best = d; bk = k; }   // minCoeff: strict <, first minimum (:21-22)
            }
        } else if (nn_mode == 1) {
            for (long k = 0; k < nm; k++) {
                double d = dist_eval(1, px, py, pz, m[3 * k], m[3 * k + 1], m[3 * k + 2]);
                if (bk < 0 || d < best) { best = d; bk = k; }
            }
        } else {
            for (long k = 0; k < nm; k++) {
                double d = dist_eval(2, px, py, pz, m[3 * k], m[3 * k + 1], m[3 * k + 2]);
                if (bk < 0 || d < best) { best = d; bk = k; }
            }
        }
        if (idx) idx[j] = (int32_t)bk;
        if (Y) { Y[3 * j] = m[3 * bk]; Y[3 * j + 1] = m[3 * bk + 1]; Y[3 * j + 2] = m[3 * bk + 2]; }  // :24
    }
}

// src/cpu.cc:29-40.  In place p_j <- sR p_j + t, returns sum_j ||Y_j - p_j||^2.  R row-major 3x3.
double oracle_err_compute(long np, double* p, double s, const double* R, const double* t, const double* Y)
{
    double err = 0.0;
    double sr[9];
    for (int i = 0; i < 9; ++i) sr[i] = s * R[i];
    for (long j = 0; j < np; j++) {
        double x = p[3 * j], y = p[3 * j + 1], z = p[3 * j + 2];
        double nx = (sr[0] * x + sr[1] * y + sr[2] * z) + t[0];
        double ny = (sr[3] * x + sr[4] * y + sr[5] * z) + t[1];
        double nz = (sr[6] * x + sr[7] * y + sr[8] * z) + t[2];
        p[3 * j] = nx; p[3 * j + 1] = ny; p[3 * j + 2] = nz;
        double ex = Y[3 * j] - nx, ey = Y[3 * j + 1] - ny, ez = Y[3 * j + 2] - nz;
        err = err + (ex * ex + ey * ey + ez * ez);
    }
    return err;
}

// src/cpu.cc:93-103.  Same sum, p not modified.
double oracle_err_compute_alignment(long np, const double* p, double s, const double* R, const double* t,
                                    const double* y)
{
    double err = 0.0;
    double sr[9];
    for (int i = 0; i < 9; ++i) sr[i] = s * R[i];
    for (long j = 0; j < np; j++) {
        double x = p[3 * j], yy = p[3 * j + 1], z = p[3 * j + 2];
        double dx = y[3 * j] - ((sr[0] * x + sr[1] * yy + sr[2] * z) + t[0]);
        double dy = y[3 * j + 1] - ((sr[3] * x + sr[4] * yy + sr[5] * z) + t[1]);
        double dz = y[3 * j + 2] - ((sr[6] * x + sr[7] * yy + sr[8] * z) + t[2]);
        err += (dx * dx + dy * dy + dz * dz);
    }
    return err;
}

// src/cpu.cc:113-117 — the "compute_centroid" step of src/bench.cc:146-157: row means and the
// centred copies.  p_prime / y_prime nullable.
void oracle_centroid(long np, const double* p, const double* y, double mu_p[3], double mu_y[3],
                     double* p_prime, double* y_prime)
{
    double sp[3] = {0, 0, 0}, sy[3] = {0, 0, 0};
    for (long j = 0; j < np; j++)
        for (int c = 0; c < 3; ++c) { sp[c] += p[3 * j + c]; sy[c] += y[3 * j + c]; }
    for (int c = 0; c < 3; ++c) { mu_p[c] = sp[c] / (double)np; mu_y[c] = sy[c] / (double)np; }
    for (long j = 0; j < np; j++)
        for (int c = 0; c < 3; ++c) {
            if (p_prime) p_prime[3 * j + c] = p[3 * j + c] - mu_p[c];
            if (y_prime) y_prime[3 * j + c] = y[3 * j + c] - mu_y[c];
        }
}

// Horn's N from the 3x3 cross-covariance S (row-major, S(a,b) = sum_j p'_a y'_b): src/cpu.cc:121-126.
void oracle_horn_matrix(const double S[9], double N[16])
{
#define s(a, b) S[(a) * 3 + (b)]
    const double n[16] = {
        s(0,0) + s(1,1) + s(2,2), s(1,2) - s(2,1), -1 * s(0,2) + s(2,0), s(0,1) - s(1,0),
        -1 * s(2,1) + s(1,2), s(0,0) - s(2,2) - s(1,1), s(0,1) + s(1,0), s(0,2) + s(2,0),
        s(2,0) - s(0,2), s(1,0) + s(0,1), s(1,1) - s(2,2) - s(0,0), s(1,2) + s(2,1),
        -1 * s(1,0) + s(0,1), s(2,0) + s(0,2), s(2,1) + s(1,2), s(2,2) - s(1,1) - s(0,0)};
#undef s
    std::memcpy(N, n, sizeof(n));
}

// Eigen-solve of N and eigenvector pick: src/cpu.cc:128-136.  Returns the picked column index in the
// solver's own eigenvalue order (eig_mode 1) or the argmax (eig_mode 0); evals_out[4] nullable.
// Returns -1 if the general solver failed.
int oracle_horn_quaternion(const double N[16], int eig_mode, double q[4], double* evals_out)
{
    if (eig_mode == 1) {
        GeneralEig4 g;
        if (!g.compute(N)) return -1;
        int k = max_element_index_literal(g.wr);
        for (int i = 0; i < 4; ++i) q[i] = g.vec[i][k];
        if (evals_out) for (int i = 0; i < 4; ++i) evals_out[i] = g.wr[i];
        return k;
    }
    double w[4], V[16];
    jacobi4(N, w, V);
    int k = 0;
    for (int i = 1; i < 4; ++i) if (w[i] > w[k]) k = i;
    for (int i = 0; i < 4; ++i) q[i] = V[i * 4 + k];
    if (evals_out) for (int i = 0; i < 4; ++i) evals_out[i] = w[i];
    return k;
}

// src/cpu.cc:138-152: R = (q_bar^T q_caps)[1:4,1:4], row-major.
// General real eigen-decomposition of a 4x4 matrix with real spectrum (Horn's N is symmetric): eigenvalues in the
// order of the quasi-triangular form's diagonal, eigenvectors (unit 2-norm) as the columns of the row-major vec.
// Used by oracle/eigen_shim/Eigen/Eigenvalues (the stand-in EigenSolver behind oracle/_ref).
int oracle_general_eig4(const double A[16], double wr[4], double wi[4], double vec[16])
{
    GeneralEig4 g;
    if (!g.compute(A)) return -1;
    for (int i = 0; i < 4; ++i) {
        wr[i] = g.wr[i];
        wi[i] = 0.0;
        for (int j = 0; j < 4; ++j) vec[i * 4 + j] = g.vec[i][j];
    }
    return 0;
}

void oracle_quat_to_rot(const double q[4], double R[9])
{
    const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
    const double q_bar[16] = {q0, -1. * q1, -1. * q2, -1. * q3,
                              q1, q0, q3, -1. * q2,
                              q2, -1. * q3, q0, q1,
                              q3, q2, -1. * q1, q0};
    const double q_caps[16] = {q0, -1. * q1, -1. * q2, -1. * q3,
                               q1, q0, -1. * q3, q2,
                               q2, q3, q0, -1. * q1,
                               q3, -1. * q2, q1, q0};
    for (int i = 1; i < 4; ++i)
        for (int j = 1; j < 4; ++j) {
            double acc = 0.0;
            for (int k = 0; k < 4; ++k) acc += q_bar[k * 4 + i] * q_caps[k * 4 + j];
            R[(i - 1) * 3 + (j - 1)] = acc;
        }
}

// src/cpu.cc:105-175.  Computes s, R (row-major), t from the current p and correspondences y and
// returns err_compute_alignment.  S_out (9, nullable) receives the cross-covariance.
double oracle_find_alignment(long np, const double* p, const double* y, int eig_mode, double* s_out,
                             double* R, double* t, double* S_out, int* picked_out)
{
    std::vector<double> pp(3 * np), yp(3 * np);
    double mu_p[3], mu_y[3];
    oracle_centroid(np, p, y, mu_p, mu_y, pp.data(), yp.data());          // :113-117
    double S[9] = {0};
    for (long j = 0; j < np; j++)                                           // :119
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) S[a * 3 + b] += pp[3 * j + a] * yp[3 * j + b];
    if (S_out) std::memcpy(S_out, S, sizeof(S));
    double N[16], q[4];
    oracle_horn_matrix(S, N);                                               // :121-126
    int picked = oracle_horn_quaternion(N, eig_mode, q, nullptr);           // :128-136
    if (picked_out) *picked_out = picked;
    oracle_quat_to_rot(q, R);                                               // :138-152
    double sp = 0., d_caps = 0.;
    for (long i = 0; i < np; i++) {                                         // :157-163
        d_caps = d_caps + ((yp[3 * i] * yp[3 * i] + yp[3 * i + 1] * yp[3 * i + 1]) + yp[3 * i + 2] * yp[3 * i + 2]);
        sp = sp + ((pp[3 * i] * pp[3 * i] + pp[3 * i + 1] * pp[3 * i + 1]) + pp[3 * i + 2] * pp[3 * i + 2]);
    }
    double s = std::sqrt(d_caps / sp);                                      // :165
    for (int a = 0; a < 3; ++a)                                             // :166-167
        t[a] = mu_y[a] - ((s * R[a * 3] * mu_p[0] + s * R[a * 3 + 1] * mu_p[1]) + s * R[a * 3 + 2] * mu_p[2]);
    *s_out = s;
    return oracle_err_compute_alignment(np, p, s, R, t, y);                 // :169-174
}

// src/cpu.cc:55-79 (+ alignement_check :42-53).  p is new_p (updated in place).
// Returns iterations performed, or -1 / -4 for the reference's two exit(-1) conditions
// (unequal sizes unless allow_unequal; fewer than 4 points).
// Per-iteration outputs (all nullable): err_hist[i] = printed "error value"; idx_hist[i*np + j];
// srt_hist[i*13 + ...] = s, R[9], t[3].  If log != 0 the reference's stderr lines are printed.
int oracle_icp_run(long nm, const double* m, long np, double* p, int max_iter, double threshold,
                   int nn_mode, int eig_mode, int allow_unequal, int nthreads, int log,
                   double* err_hist, int32_t* idx_hist, double* srt_hist)
{
    if (np != nm && !allow_unequal) {
        if (log) std::fprintf(stderr, "[error] Point sets need to have the same number of points.\n");
        return -1;
    }
    if (np < 4) {
        if (log) std::fprintf(stderr, "[error] Need at least 4 point pairs\n");
        return -4;
    }
    std::vector<double> Y(3 * np);
    std::vector<int32_t> idx(np);
    int it = 0;
    for (int i = 0; i < max_iter; i++) {
        if (log) std::fprintf(stderr, "[ICP] iteration number %d | ", i);
        oracle_closest_matrix(np, nm, p, m, nn_mode, nthreads, idx.data(), Y.data());   // :63
        double s, R[9], t[3];
        double err = oracle_find_alignment(np, p, Y.data(), eig_mode, &s, R, t, nullptr, nullptr);  // :65
        err += oracle_err_compute(np, p, s, R, t, Y.data());                             // :67-71
        err /= (double)np;                                                               // :73
        if (log) { std::ostringstream os; os << "error value = " << err << std::endl; std::fputs(os.str().c_str(), stderr); }
        if (err_hist) err_hist[i] = err;
        if (idx_hist) std::memcpy(idx_hist + (size_t)i * np, idx.data(), sizeof(int32_t) * np);
        if (srt_hist) { srt_hist[i * 13] = s; std::memcpy(srt_hist + i * 13 + 1, R, 72); std::memcpy(srt_hist + i * 13 + 10, t, 24); }
        it = i + 1;
        if (err < threshold) break;                                                      // :76-77
    }
    return it;
}

}  // extern "C"
