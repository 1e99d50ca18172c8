// icp.hh — C++ host mirror of the reference's ICP interface over the C-ABI (include/icp_b200.h).
//
// A maintainer of yassram/iterative-closest-point can swap `#include "cpu.hh"` / `"GPU/gpu.hh"` for this
// header: the names, argument meaning and error behaviour of the hot-path entry points are kept
//   CPU::closest_matrix / err_compute / err_compute_alignment        src/cpu.hh:116-118
//   CPU::ICP {ctor, find_corresponding, alignement_check, find_alignment, new_p}   src/cpu.hh:45-114
//   GPU::ICP {find_corresponding_opti, find_corresponding_naive, compute_y_naive, getDim, getNp}  src/GPU/gpu.hh:41-104
//   compute_Y_w_opti, compute_distance_w_naive, compute_err_w, substract_col_w, y_p_norm_w   src/GPU/gpu.hh:110-116
// while every body is one call into libicp_b200.so (CUDA, sm_100a).  No arithmetic happens here
// and there is no CPU fallback.  Eigen is not required: icpb::Matrix is a minimal column-major
// double matrix whose memory image equals Eigen::MatrixXd's, so `Eigen::Map` adapts it for free.
#pragma once
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <vector>

#include "icp_b200.h"

namespace icpb {

class Matrix {
public:
    Matrix() = default;
    Matrix(long rows, long cols) : r_(rows), c_(cols), d_((size_t)(rows * cols), 0.0) {}
    static Matrix Zero(long rows, long cols) { return Matrix(rows, cols); }
    static Matrix Identity(long rows, long cols)
    {
        Matrix m(rows, cols);
        for (long i = 0; i < rows && i < cols; ++i) m(i, i) = 1.0;
        return m;
    }
    long rows() const { return r_; }
    long cols() const { return c_; }
    double* data() { return d_.data(); }
    const double* data() const { return d_.data(); }
    double& operator()(long i, long j) { return d_[(size_t)(j * r_ + i)]; }
    double operator()(long i, long j) const { return d_[(size_t)(j * r_ + i)]; }
    double& operator()(long i) { return d_[(size_t)i]; }
    double operator()(long i) const { return d_[(size_t)i]; }

#ifdef ICP_B200_EIGEN_STANDIN
    // host/compat only: the two Eigen expressions the reference's benchmark harness evaluates itself
    // (`M.rowwise().mean()`, `M.colwise() - v`, src/bench.cc:150-154,164-165).  They are the CALLER's Eigen code, not
    // entry points of the path; with the real Eigen present this block is not compiled (see compat/cpu.hh).
    struct RowwiseProxy {
        const Matrix& m;
        Matrix mean() const
        {
            Matrix out(m.rows(), 1);
            for (long j = 0; j < m.cols(); ++j)
                for (long i = 0; i < m.rows(); ++i) out(i) += m(i, j);
            for (long i = 0; i < m.rows(); ++i) out(i) /= (double)m.cols();
            return out;
        }
    };
    struct ColwiseProxy {
        const Matrix& m;
        Matrix operator-(const Matrix& v) const
        {
            Matrix out(m.rows(), m.cols());
            for (long j = 0; j < m.cols(); ++j)
                for (long i = 0; i < m.rows(); ++i) out(i, j) = m(i, j) - v(i);
            return out;
        }
    };
    RowwiseProxy rowwise() const { return {*this}; }
    ColwiseProxy colwise() const { return {*this}; }
#endif

private:
    long r_ = 0, c_ = 0;
    std::vector<double> d_;
};

[[noreturn]] inline void die(const char* where, int rc)
{
    std::cerr << "[icp_b200] " << where << " failed (" << rc << "): " << icp_last_error() << std::endl;
    std::exit(3);
}

// Device contexts of ICP objects are recycled per host thread: creating one costs ~3 ms (stream, events, a dozen
// cudaMalloc), several times a whole cow-sized loop, and the reference's harness constructs an ICP per repetition
// (src/bench.cc:71-105).  A recycled context keeps its (grow-only) device buffers; icp_set_model / icp_set_scene reset
// everything else.  Contexts still pooled at exit are left to process teardown.
inline std::vector<icp_ctx*>& ctx_pool()
{
    static thread_local std::vector<icp_ctx*> pool;
    return pool;
}
inline int acquire_ctx(icp_ctx** out)
{
    auto& pool = ctx_pool();
    if (!pool.empty()) {
        *out = pool.back();
        pool.pop_back();
        return ICP_OK;
    }
    return icp_create(out, 0);
}
inline void release_ctx(icp_ctx* c)
{
    if (!c) return;
    if (ctx_pool().size() < 4) ctx_pool().push_back(c);
    else icp_destroy(c);
}

// ICP_DEVICES=0,1,...  (environment): ICP objects run their loop on those GPUs of this process — the scene cut into
// contiguous blocks, the model replicated, the exchange inside the alignment kernels over peer memory (icp_group_*,
// include/icp_b200.h).  Unset or a single device: one GPU, as before.  Results do not depend on it.
inline std::vector<int> env_devices()
{
    std::vector<int> d;
    if (const char* e = std::getenv("ICP_DEVICES")) {
        for (const char* q = e; *q;) {
            char* end = nullptr;
            const long v = std::strtol(q, &end, 10);
            if (end == q) break;
            d.push_back((int)v);
            q = (*end == ',') ? end + 1 : end;
        }
    }
    return d;
}

// row-major double[9] <-> column-major 3x3 Matrix
inline void to_rowmajor3(const Matrix& m, double out[9])
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) out[i * 3 + j] = m(i, j);
}
inline Matrix from_rowmajor3(const double in[9])
{
    Matrix m(3, 3);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) m(i, j) = in[i * 3 + j];
    return m;
}

struct closest_matrix_params {
    unsigned int dim, np, nm;
    const Matrix& p;
    const Matrix& m;
};
struct err_compute_params {
    unsigned int np;
    Matrix& p;
    double s;
    const Matrix& r;
    const Matrix& t;
    const Matrix& Y;
};
struct err_compute_alignment_params {
    unsigned int np;
    const Matrix& p;
    double s;
    const Matrix& r;
    const Matrix& t;
    const Matrix& y;
};

// src/GPU/gpu.hh:25-39
struct gpu_closest_matrix_params {
    const Matrix& p;
    const Matrix& m;
    Matrix& y;
};
struct gpu_err_compute_params {
    const Matrix& Y;
    Matrix& p;
    bool s;               // the in_place flag of compute_err_w (the reference names it `s`)
    const Matrix sr;      // s * R
    const Matrix& t;
};

// src/cpu.cc:5-27
inline Matrix closest_matrix(closest_matrix_params pa)
{
    Matrix res(pa.dim, pa.np);
    int rc = icp_closest_matrix_host(pa.m.data(), pa.nm, pa.p.data(), pa.np, nullptr, res.data());
    if (rc) die("closest_matrix", rc);
    return res;
}

inline double err_common(const Matrix& Y, Matrix& p, long np, bool in_place, double s, const Matrix& r, const Matrix& t)
{
    double sr[9], tt[3] = {t(0), t(1), t(2)}, err = 0.0;
    to_rowmajor3(r, sr);
    for (double& v : sr) v *= s;
    int rc = icp_err_compute_host(Y.data(), p.data(), np, in_place ? 1 : 0, sr, tt, &err);
    if (rc) die("err_compute", rc);
    return err;
}
// src/cpu.cc:29-40: p is updated in place through the reference member
inline double err_compute(err_compute_params pa) { return err_common(pa.Y, pa.p, pa.np, true, pa.s, pa.r, pa.t); }
// src/cpu.cc:93-103
inline double err_compute_alignment(err_compute_alignment_params pa)
{
    return err_common(pa.y, const_cast<Matrix&>(pa.p), pa.np, false, pa.s, pa.r, pa.t);
}

class ICP {
public:
    ICP(Matrix m_, Matrix p_, int max_iter_)
        : new_p(p_), m(std::move(m_)), p(std::move(p_)), np((unsigned)p.cols()), nm((unsigned)m.cols()), dim(3),
          max_iter(max_iter_)
    {
        s = 1.;
        r = Matrix::Identity(3, 3);
        t = Matrix::Zero(3, 1);
        ensure_ctx();
    }
    // The reference's ICP objects are plain host values and its harness copies them freely (src/bench.cc:186-189,
    // 211-226, 362-381).  A copy takes the host state (m, p, new_p, s, r, t); its device context is created on
    // first use from the copy's current new_p.
    ICP(const ICP& o)
        : new_p(o.new_p), iterations(o.iterations), errors(o.errors), s(o.s), t(o.t), r(o.r), m(o.m), p(o.p), np(o.np),
          nm(o.nm), dim(o.dim), max_iter(o.max_iter)
    {
    }
    ICP& operator=(const ICP&) = delete;
    ~ICP()
    {
        release_ctx(ctx);
        if (grp) icp_group_destroy(grp);
    }

    closest_matrix_params get_closest_matrix_params() { return {dim, np, nm, new_p, m}; }
    err_compute_params get_err_compute_params(Matrix& Y) { return {np, new_p, s, r, t, Y}; }
    err_compute_alignment_params get_err_compute_alignment_params(Matrix& Y) { return {np, new_p, s, r, t, Y}; }
    // GPU::ICP's forms, src/GPU/gpu.hh:57-72
    gpu_closest_matrix_params get_closest_matrix_params(Matrix& Y) { return {new_p, m, Y}; }
    gpu_err_compute_params get_err_compute_params(Matrix& Y, bool in_place)
    {
        Matrix sr(3, 3);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) sr(i, j) = s * r(i, j);
        return {Y, new_p, in_place, sr, t};
    }
    double getDim() { return dim; }
    double getNp() { return np; }

    // src/cpu.cc:42-53
    void alignement_check()
    {
        if (p.cols() != m.cols()) {
            std::cerr << "[error] Point sets need to have the same number of points.\n";
            std::exit(-1);
        }
        if (p.cols() < 4) {
            std::cerr << "[error] Need at least 4 point pairs\n";
            std::exit(-1);
        }
    }

    // src/cpu.cc:55-79 — the whole loop runs on the device; the reference's per-iteration stderr lines
    // are emitted afterwards from the recorded error values.
    void find_corresponding(bool check_sizes = true)
    {
        if (check_sizes) alignement_check();
        ensure_ctx();
        std::vector<double> hist((size_t)(max_iter > 0 ? max_iter : 1));
        int rc = grp ? icp_group_run(grp, max_iter, threshold, 0, &iterations, hist.data())
                     : icp_run(ctx, max_iter, threshold, 0, &iterations, hist.data());
        if (rc) die("icp_run", rc);
        for (int i = 0; i < iterations; i++) {
            std::cerr << "[ICP] iteration number " << i << " | ";
            std::cerr << "error value = " << hist[(size_t)i] << std::endl;
        }
        errors.assign(hist.begin(), hist.begin() + iterations);
        if ((rc = grp ? icp_group_get_scene(grp, new_p.data()) : icp_get_scene(ctx, new_p.data()))) die("icp_get_scene", rc);
        pull_transform();
    }
    void find_corresponding_opti() { find_corresponding(); }    // src/GPU/gpu.cc:52-83
    void find_corresponding_naive() { find_corresponding(); }   // src/GPU/gpu.cc:16-49 (same results)

    Matrix compute_y_naive()                                     // src/GPU/gpu.cc:6-14
    {
        Matrix Y(dim, np);
        ensure_ctx();
        int rc = grp ? icp_group_closest(grp, nullptr, Y.data()) : icp_closest(ctx, nullptr, Y.data());
        if (rc) die("icp_closest", rc);
        return Y;
    }

    // src/cpu.cc:105-175: s, R, t from new_p and the given correspondences y; returns err_compute_alignment.
    double find_alignment(const Matrix& y)
    {
        double R_[9], t_[3], err = 0.0;
        int rc = icp_find_alignment_host(y.data(), new_p.data(), np, &s, R_, t_, &err);
        if (rc) die("icp_find_alignment_host", rc);
        r = from_rowmajor3(R_);
        for (int i = 0; i < 3; ++i) t(i) = t_[i];
        return err;
    }

    double scale() const { return s; }
    const Matrix& rotation() const { return r; }
    const Matrix& translation() const { return t; }
    icp_ctx* handle() { ensure_ctx(); return grp ? icp_group_member(grp, 0) : ctx; }
    icp_group* group() { ensure_ctx(); return grp; }

public:
    Matrix new_p;
    int iterations = 0;
    std::vector<double> errors;

private:
    void ensure_ctx()     // the constructor, and copies on first use
    {
        if (ctx || grp) return;
        const std::vector<int> devs = env_devices();
        int rc;
        if (devs.size() > 1 && np >= devs.size()) {
            if ((rc = icp_group_create(&grp, devs.data(), (int)devs.size()))) die("icp_group_create", rc);
            if ((rc = icp_group_set_model(grp, m.data(), nm))) die("icp_group_set_model", rc);
            if ((rc = icp_group_set_scene(grp, new_p.data(), np))) die("icp_group_set_scene", rc);
            return;
        }
        if (devs.size() == 1) rc = icp_create(&ctx, devs[0]);
        else rc = acquire_ctx(&ctx);
        if (rc) die("icp_create", rc);
        if ((rc = icp_set_model(ctx, m.data(), nm))) die("icp_set_model", rc);
        if ((rc = icp_set_scene(ctx, new_p.data(), np))) die("icp_set_scene", rc);
    }
    void pull_transform()
    {
        double R_[9], t_[3];
        int rc = grp ? icp_group_get_transform(grp, &s, R_, t_) : icp_get_transform(ctx, &s, R_, t_);
        if (rc) die("icp_get_transform", rc);
        r = from_rowmajor3(R_);
        for (int i = 0; i < 3; ++i) t(i) = t_[i];
    }
    double s;
    Matrix t, r, m, p;
    unsigned int np, nm, dim;
    int max_iter;
    const double threshold = 1e-5;   // src/cpu.hh:113
    icp_ctx* ctx = nullptr;
    icp_group* grp = nullptr;        // ICP_DEVICES lists several GPUs
};

}  // namespace icpb

// The reference exposes the same path under two namespaces (CPU:: in src/cpu.hh, GPU:: in src/GPU/gpu.hh).
namespace CPU {
using icpb::closest_matrix;
using icpb::closest_matrix_params;
using icpb::err_compute;
using icpb::err_compute_alignment;
using icpb::err_compute_alignment_params;
using icpb::err_compute_params;
using icpb::ICP;
}  // namespace CPU
namespace GPU {
using icpb::gpu_closest_matrix_params;
using icpb::gpu_err_compute_params;
using icpb::ICP;
using icpb::Matrix;
}  // namespace GPU

// ---- operator-level wrappers, src/GPU/gpu.hh:110-116 ---------------------------------------------
inline void compute_Y_w_opti(const GPU::Matrix& m, const GPU::Matrix& pi, GPU::Matrix& Y)
{
    if (Y.rows() != 3 || Y.cols() != pi.cols()) Y = GPU::Matrix(3, pi.cols());
    int rc = icp_closest_matrix_host(m.data(), m.cols(), pi.data(), pi.cols(), nullptr, Y.data());
    if (rc) icpb::die("compute_Y_w_opti", rc);
}
inline int compute_distance_w_naive(const GPU::Matrix& m, const GPU::Matrix& pi)
{
    int32_t k = -1;
    int rc = icp_nearest_index_host(m.data(), m.cols(), pi.data(), &k);
    if (rc) icpb::die("compute_distance_w_naive", rc);
    return k;
}
inline double compute_err_w(const GPU::Matrix& Y, GPU::Matrix& p, bool in_place, const GPU::Matrix& sr,
                            const GPU::Matrix& t)
{
    double srr[9], tt[3] = {t(0), t(1), t(2)}, err = 0.0;
    icpb::to_rowmajor3(sr, srr);
    int rc = icp_err_compute_host(Y.data(), p.data(), p.cols(), in_place ? 1 : 0, srr, tt, &err);
    if (rc) icpb::die("compute_err_w", rc);
    return err;
}
inline GPU::Matrix substract_col_w(const GPU::Matrix& M, const GPU::Matrix& m)
{
    GPU::Matrix out(M.rows(), M.cols());
    const double mm[3] = {m(0), m(1), m(2)};
    int rc = icp_subtract_col_host(M.data(), M.cols(), mm, out.data());
    if (rc) icpb::die("substract_col_w", rc);
    return out;
}
inline void y_p_norm_w(const GPU::Matrix& y, const GPU::Matrix& p, size_t size_arr, double& d_caps, double& sp)
{
    int rc = icp_y_p_norm_host(y.data(), p.data(), (int64_t)size_arr, &d_caps, &sp);
    if (rc) icpb::die("y_p_norm_w", rc);
}
