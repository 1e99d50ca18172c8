// ref_wrap.cc — C entry points over the REFERENCE'S OWN src/cpu.cc (compiled unmodified, in place, from
// /root/reference/src by oracle/Makefile's `_ref` target against the Eigen stand-in of oracle/eigen_shim).
// TEST INFRASTRUCTURE ONLY: used by tests/test_ref_cpu.py to check the oracle restatement against the
// reference's actual source text.  Clouds cross as n x 3 doubles (x0 y0 z0 x1 ...), the memory image of the
// reference's column-major 3 x n MatrixXd.
#include <cstring>

#include "cpu.hh"      // /root/reference/src/cpu.hh

namespace {
MatrixXd to_mat(const double* xyz, long n)
{
    MatrixXd m(3, n);
    std::memcpy(m.data(), xyz, (size_t)n * 24);
    return m;
}
}  // namespace

extern "C" {

// CPU::closest_matrix (src/cpu.cc:5-27): Y out (np x 3).
void ref_closest_matrix(long np, long nm, const double* p, const double* m, double* Y_out)
{
    MatrixXd P = to_mat(p, np), M = to_mat(m, nm);
    CPU::closest_matrix_params pa{3u, (unsigned)np, (unsigned)nm, P, M};
    MatrixXd Y = CPU::closest_matrix(pa);
    std::memcpy(Y_out, Y.data(), (size_t)np * 24);
}

// CPU::err_compute (src/cpu.cc:29-40): p updated in place.
double ref_err_compute(long np, double* p, double s, const double* R_rowmajor, const double* t, const double* Y)
{
    MatrixXd P = to_mat(p, np), Ym = to_mat(Y, np), R(3, 3), T(3, 1);
    for (int i = 0; i < 3; ++i) {
        T(i, 0) = t[i];
        for (int j = 0; j < 3; ++j) R(i, j) = R_rowmajor[i * 3 + j];
    }
    CPU::err_compute_params pa{(unsigned)np, P, s, R, T, Ym};
    const double e = CPU::err_compute(pa);
    std::memcpy(p, P.data(), (size_t)np * 24);
    return e;
}

// CPU::err_compute_alignment (src/cpu.cc:93-103).
double ref_err_compute_alignment(long np, const double* p, double s, const double* R_rowmajor, const double* t,
                                 const double* y)
{
    MatrixXd P = to_mat(p, np), Ym = to_mat(y, np), R(3, 3), T(3, 1);
    for (int i = 0; i < 3; ++i) {
        T(i, 0) = t[i];
        for (int j = 0; j < 3; ++j) R(i, j) = R_rowmajor[i * 3 + j];
    }
    CPU::err_compute_alignment_params pa{(unsigned)np, P, s, R, T, Ym};
    return CPU::err_compute_alignment(pa);
}

// CPU::ICP::find_alignment (src/cpu.cc:105-175) on new_p = p and correspondences y; the private s, r, t are
// recovered through get_err_compute_params (src/cpu.hh:75-82), which exposes references to them.
double ref_find_alignment(long np, const double* p, const double* y, double* s_out, double* R_rowmajor_out, double* t_out)
{
    MatrixXd P = to_mat(p, np), Ym = to_mat(y, np);
    CPU::ICP icp(P, P, 1);
    const double err = icp.find_alignment(Ym);
    CPU::err_compute_params ecp = icp.get_err_compute_params(Ym);
    *s_out = ecp.s;
    for (int i = 0; i < 3; ++i) {
        t_out[i] = ecp.t(i, 0);
        for (int j = 0; j < 3; ++j) R_rowmajor_out[i * 3 + j] = ecp.r(i, j);
    }
    return err;
}

// CPU::ICP ctor + find_corresponding (src/cpu.cc:55-79): returns nothing (the reference prints its error values
// on stderr); new_p out.  exit(-1) on unequal sizes, as the reference does.
void ref_find_corresponding(long nm, const double* m, long np, const double* p, int max_iter, double* new_p_out)
{
    CPU::ICP icp(to_mat(m, nm), to_mat(p, np), max_iter);
    icp.find_corresponding();
    std::memcpy(new_p_out, icp.new_p.data(), (size_t)np * 24);
}

}  // extern "C"
