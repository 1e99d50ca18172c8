// bench.cc — the reference's benchmark list (src/bench.cc:391-445, google-benchmark captures on
// data_students/cow_ref.txt / cow_tr1.txt, UseRealTime, milliseconds, max_iter 20 for the loops) re-stated
// on the B200 path without google-benchmark.  One line per reference capture, under the reference's own
// capture name; the CPU:: and GPU:: captures of the reference both map onto the same B200 operators (the
// drop-in has one implementation), so each operator is timed
//   one-shot   through the reference-shaped stateless entry point (host matrices in, host results out: context
//              creation, uploads, Morton sort, kernels, downloads — what BM_GPU_* measures in the reference)
//   resident   on a context that already holds the clouds (kernels + result read-back only)
// Usage: bench [ref.txt scene.txt] [--json] [--devices 0,1,...]     (defaults: the reference's REF_PATH / SCENE_PATH)
//   --devices  adds the resident loop over several GPUs of this process (icp_group_*: the scene cut into contiguous
//              blocks, strong scaling) next to the single-GPU lines; ICP_DEVICES in the environment moves the ICP objects
//              of the one-shot lines too (host/icp.hh).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "load.hh"

namespace {

using clk = std::chrono::steady_clock;

struct Result {
    std::string name, maps_to;
    double mean_ms, min_ms;
    int reps;
};

// google-benchmark-like: warm up, then repeat until >= 0.5 s of samples or 200 repetitions (at least 5)
Result measure(const std::string& name, const std::string& maps_to, const std::function<void()>& fn)
{
    for (int i = 0; i < 3; ++i) fn();
    std::vector<double> t;
    double total = 0.0;
    while ((total < 500.0 && t.size() < 200) || t.size() < 5) {
        const auto t0 = clk::now();
        fn();
        const double ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
        t.push_back(ms);
        total += ms;
    }
    return {name, maps_to, total / (double)t.size(), *std::min_element(t.begin(), t.end()), (int)t.size()};
}

#define CHECK(call) do { int rc__ = (call); if (rc__) icpb::die(#call, rc__); } while (0)

}  // namespace

int main(int argc, char* argv[])
{
    const char* ref_path = "data_students/cow_ref.txt";
    const char* scene_path = "data_students/cow_tr1.txt";
    bool json = false;
    std::vector<const char*> pos;
    std::vector<int> devices;
    for (int i = 1; i < argc; ++i) {
        if (!std::strcmp(argv[i], "--json")) json = true;
        else if (!std::strcmp(argv[i], "--devices") && i + 1 < argc) {
            for (const char* q = argv[++i]; *q;) {
                char* end = nullptr;
                const long v = std::strtol(q, &end, 10);
                if (end == q) break;
                devices.push_back((int)v);
                q = (*end == ',') ? end + 1 : end;
            }
        } else pos.push_back(argv[i]);
    }
    if (pos.size() >= 2) { ref_path = pos[0]; scene_path = pos[1]; }
    const int loop_iters = 20;                                      // src/bench.cc:385,435-445

    GPU::Matrix m = load_matrix(ref_path), p = load_matrix(scene_path);
    const long np = p.cols(), nm = m.cols();
    std::vector<Result> out;

    // ---- one-shot forms (reference-shaped wrappers) ---------------------------------------------------
    GPU::Matrix Y(3, np);
    out.push_back(measure("opti_gpu_closest_matrix", "BM_GPU_closest_matrix_opti / BM_CPU_closest_matrix (one-shot)",
                          [&] { compute_Y_w_opti(m, p, Y); }));
    out.push_back(measure("naive_gpu_closest_matrix", "BM_GPU_closest_matrix_naive: one scene point per call (x8 calls)", [&] {
        GPU::Matrix pi(3, 1);
        for (int j = 0; j < 8; ++j) {
            for (int c = 0; c < 3; ++c) pi(c) = p(c, j);
            (void)compute_distance_w_naive(m, pi);
        }
    }));
    {
        CPU::ICP icp(m, p, loop_iters);
        out.push_back(measure("gpu_find_alignment", "BM_GPU_find_alignment / BM_CPU_find_alignment (one-shot)",
                              [&] { (void)icp.find_alignment(Y); }));
        GPU::Matrix mu(3, 1);
        out.push_back(measure("gpu_compute_centroid", "BM_GPU_Compute_centroid / BM_CPU_Compute_centroid (substract_col_w x2, one-shot)",
                              [&] { (void)substract_col_w(p, mu); (void)substract_col_w(Y, mu); }));
        GPU::Matrix sr = GPU::Matrix::Identity(3, 3), t = GPU::Matrix::Zero(3, 1), pc = p;
        out.push_back(measure("gpu_err_compute", "BM_GPU_Err_compute(in_place) / BM_CPU_Err_compute (one-shot)",
                              [&] { (void)compute_err_w(Y, pc, true, sr, t); }));
        out.push_back(measure("gpu_err_compute_alignment", "BM_GPU_Err_compute(not in place) / BM_CPU_Err_compute_alignment (one-shot)",
                              [&] { (void)compute_err_w(Y, pc, false, sr, t); }));
    }
    out.push_back(measure("opti_gpu_loop", "BM_GPU_Find_corresponding_opti / _naive / BM_CPU_Find_corresponding (ctor + loop, max_iter 20)", [&] {
        CPU::ICP icp(m, p, loop_iters);
        icp.find_corresponding(np == nm);
    }));

    // ---- resident forms (clouds already in HBM) -------------------------------------------------------
    icp_ctx* ctx = nullptr;
    CHECK(icp_create(&ctx, 0));
    CHECK(icp_set_model(ctx, m.data(), nm));
    CHECK(icp_set_scene(ctx, p.data(), np));
    out.push_back(measure("resident_closest_matrix", "icp_closest (a1), no host output", [&] { CHECK(icp_closest(ctx, nullptr, nullptr)); }));
    double mu_p[3], mu_y[3], s, R[9], t3[3], err;
    out.push_back(measure("resident_compute_centroid", "icp_centroid (a2)", [&] { CHECK(icp_centroid(ctx, mu_p, mu_y)); }));
    out.push_back(measure("resident_find_alignment", "icp_find_alignment (a2-a7)", [&] { CHECK(icp_find_alignment(ctx, &s, R, t3, &err)); }));
    out.push_back(measure("resident_err_compute_alignment", "icp_err_compute(in_place=0) (a6)", [&] { CHECK(icp_err_compute(ctx, 0, &err)); }));
    out.push_back(measure("resident_err_compute", "icp_err_compute(in_place=1) (a8)", [&] { CHECK(icp_err_compute(ctx, 1, &err)); }));
    int iters = 0;
    std::vector<double> hist((size_t)loop_iters);
    out.push_back(measure("resident_loop", "icp_reset_scene + icp_run(max_iter 20) (a9)", [&] {
        CHECK(icp_reset_scene(ctx));
        CHECK(icp_run(ctx, loop_iters, 1e-5, 0, &iters, hist.data()));
    }));
    icp_stats st;
    CHECK(icp_get_stats(ctx, &st));
    icp_destroy(ctx);

    // ---- the same resident loop over several GPUs of this process -------------------------------------
    icp_stats gst{};
    int giters = 0;
    if (devices.size() > 1) {
        icp_group* grp = nullptr;
        CHECK(icp_group_create(&grp, devices.data(), (int)devices.size()));
        CHECK(icp_group_set_model(grp, m.data(), nm));
        CHECK(icp_group_set_scene(grp, p.data(), np));
        out.push_back(measure("resident_loop_group", "icp_group_reset_scene + icp_group_run(max_iter 20) on " + std::to_string(devices.size()) + " GPUs (a9, scene sharded)", [&] {
            CHECK(icp_group_reset_scene(grp));
            CHECK(icp_group_run(grp, loop_iters, 1e-5, 0, &giters, hist.data()));
        }));
        CHECK(icp_group_get_stats(grp, &gst));
        icp_group_destroy(grp);
        if (giters != iters) { std::fprintf(stderr, "[bench] the group loop ran %d iterations, the single-GPU loop %d\n", giters, iters); return 4; }
    }

    if (json) {
        std::printf("{\"ref\": \"%s\", \"scene\": \"%s\", \"np\": %ld, \"nm\": %ld, \"loop_iterations\": %d, \"benchmarks\": [", ref_path,
                    scene_path, np, nm, iters);
        for (size_t i = 0; i < out.size(); ++i)
            std::printf("%s{\"name\": \"%s\", \"real_time_ms\": %.6f, \"min_ms\": %.6f, \"repetitions\": %d, \"frame_rate\": %.3f}",
                        i ? ", " : "", out[i].name.c_str(), out[i].mean_ms, out[i].min_ms, out[i].reps, 1000.0 / out[i].mean_ms);
        std::printf("]}\n");
    } else {
        std::printf("%s vs %s: np = %ld, nm = %ld; loops stop after %d iterations; device loop %.3f ms\n", ref_path, scene_path, np, nm,
                    iters, st.ms_total);
        if (devices.size() > 1)
            std::printf("group of %zu GPUs: %d iterations, device loop %.3f ms (slowest device), filter %.3f ms\n", devices.size(), giters,
                        gst.ms_total, gst.ms_filter);
        std::printf("%-34s %12s %12s %6s %12s   %s\n", "Benchmark", "Time (ms)", "min (ms)", "reps", "frame_rate", "reference capture");
        for (const auto& r : out)
            std::printf("%-34s %12.4f %12.4f %6d %12.2f   %s\n", r.name.c_str(), r.mean_ms, r.min_ms, r.reps, 1000.0 / r.mean_ms,
                        r.maps_to.c_str());
    }
    return 0;
}
