// compat/benchmark/benchmark.h — the slice of google-benchmark's interface that the reference's harness uses
// (src/bench.cc: BENCHMARK_CAPTURE(...)->Unit(kMillisecond)->UseRealTime(), `for (auto _ : state)`,
// state.iterations(), state.counters[...] = Counter(n, Counter::kIsRate), BENCHMARK_MAIN()), so that src/bench.cc
// builds without that dependency.  Timing rule as in google-benchmark: the captured function is re-run with a growing
// iteration count until one run lasts at least min_time (0.5 s; ICP_BENCH_MIN_TIME overrides), real time, and the
// per-iteration mean of that last run is printed.  The capture's arguments are re-evaluated on every run, as there.
#pragma once
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <vector>

namespace benchmark {

enum TimeUnit { kNanosecond, kMicrosecond, kMillisecond, kSecond };

struct Counter {
    enum Flags { kDefaults = 0, kIsRate = 1 };
    double value = 0.0;
    int flags = kDefaults;
    Counter() = default;
    Counter(double v, int f = kDefaults) : value(v), flags(f) {}
};

class State {
public:
    explicit State(long long n) : max_(n) {}
    struct Token {};
    struct Iterator {
        State* st;
        long long left;
        bool operator!=(const Iterator&) {
            if (left > 0) return true;
            st->finish();
            return false;
        }
        Iterator& operator++() { --left; return *this; }
        Token operator*() const { return {}; }
    };
    Iterator begin() { start_ = std::chrono::steady_clock::now(); started_ = true; return {this, max_}; }
    Iterator end() { return {this, 0}; }
    long long iterations() const { return max_; }
    std::map<std::string, Counter> counters;
    double seconds() const { return sec_; }

private:
    void finish() { if (started_) { sec_ = std::chrono::duration<double>(std::chrono::steady_clock::now() - start_).count(); started_ = false; } }
    long long max_;
    std::chrono::steady_clock::time_point start_{};
    bool started_ = false;
    double sec_ = 0.0;
};

namespace internal {

struct Benchmark {
    std::string name;
    std::function<void(State&)> fn;
    TimeUnit unit = kNanosecond;
    Benchmark* Unit(TimeUnit u) { unit = u; return this; }
    Benchmark* UseRealTime() { return this; }   // real time is the only clock here
};

inline std::vector<Benchmark*>& registry()
{
    static std::vector<Benchmark*> r;
    return r;
}
inline Benchmark* Register(const char* name, std::function<void(State&)> fn)
{
    Benchmark* b = new Benchmark{name, std::move(fn)};
    registry().push_back(b);
    return b;
}

inline int RunAll(int argc, char** argv)
{
    double min_time = 0.5;
    if (const char* e = std::getenv("ICP_BENCH_MIN_TIME")) min_time = std::atof(e);
    const char* filter = nullptr;
    for (int i = 1; i < argc; ++i)
        if (!std::strncmp(argv[i], "--benchmark_filter=", 19)) filter = argv[i] + 19;
    static const char* unit_name[] = {"ns", "us", "ms", "s"};
    static const double unit_scale[] = {1e9, 1e6, 1e3, 1.0};
    std::printf("%-55s %15s %12s %s\n", "Benchmark", "Time", "Iterations", "UserCounters...");
    for (Benchmark* b : registry()) {
        if (filter && !std::strstr(b->name.c_str(), filter)) continue;
        long long n = 1;
        double sec = 0.0;
        double rate = 0.0;
        for (;;) {
            State st(n);
            b->fn(st);
            sec = st.seconds();
            for (auto& kv : st.counters)
                if (kv.first == "frame_rate") rate = (kv.second.flags & Counter::kIsRate) && sec > 0 ? kv.second.value / sec : kv.second.value;
            if (sec >= min_time || n >= 1000000000LL) break;
            double mult = sec > 0 ? min_time * 1.4 / sec : 10.0;
            if (mult > 10.0 || sec < min_time * 0.1) mult = 10.0;
            long long next = (long long)((double)n * mult) + 1;
            n = next > n ? next : n + 1;
        }
        std::printf("%-55s %12.3f %s %12lld frame_rate=%g/s\n", (b->name + "/real_time").c_str(),
                    sec / (double)n * unit_scale[b->unit], unit_name[b->unit], n, rate);
        std::fflush(stdout);
    }
    return 0;
}

}  // namespace internal
}  // namespace benchmark

#define ICP_BM_CAT2(a, b) a##b
#define ICP_BM_CAT(a, b) ICP_BM_CAT2(a, b)
#define BENCHMARK_CAPTURE(func, test_case_name, ...)                                                        \
    static ::benchmark::internal::Benchmark* ICP_BM_CAT(icp_bm_, __LINE__) = ::benchmark::internal::Register( \
        #func "/" #test_case_name, [](::benchmark::State& st) { func(st, __VA_ARGS__); })
#define BENCHMARK_MAIN() \
    int main(int argc, char** argv) { return ::benchmark::internal::RunAll(argc, argv); }
