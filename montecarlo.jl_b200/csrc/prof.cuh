// Launch accounting: every kernel launch of the library is counted per class; when profiling is
// switched on for a handle each launch is also bracketed by a pair of CUDA events on the handle's
// stream, so bench.py can report the dominant kernel's live average duration (roofline.achieved).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

namespace mcb {

enum LaunchClass { LC_GEMM = 0, LC_UDT, LC_RDIVP, LC_SWEEP, LC_FUSED, LC_MEASURE, LC_MISC, LC_ISING, LC_ADVANCE, LC_GREENS, LC_COUNT };

struct Profiler {
    bool timing = false;
    int64_t launches[LC_COUNT] = {};
    double flops[LC_COUNT] = {};   // algorithmic FP64 flop handed to the launches of each class
    struct Span { int cls; cudaEvent_t a, b; };
    std::vector<Span> spans;
    double ms[LC_COUNT] = {};
    int64_t timed[LC_COUNT] = {};

    void reset();
    // resolves outstanding event pairs into ms[] (the stream must have been synchronised)
    void collect();
};

// the profiler launches are currently attributed to (set by the C-ABI entry points)
void set_current_profiler(Profiler* p);
Profiler* current_profiler();

// RAII bracket around one launch
struct LaunchScope {
    Profiler* p;
    cudaStream_t st;
    int idx = -1;
    LaunchScope(int cls, cudaStream_t st, double flop);
    ~LaunchScope();
};

} // namespace mcb
