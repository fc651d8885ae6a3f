#include "prof.cuh"

namespace mcb {

static thread_local Profiler* g_prof = nullptr;
void set_current_profiler(Profiler* p) { g_prof = p; }
Profiler* current_profiler() { return g_prof; }

void Profiler::reset()
{
    collect();
    for (int i = 0; i < LC_COUNT; ++i) { launches[i] = 0; flops[i] = 0; ms[i] = 0; timed[i] = 0; }
}

void Profiler::collect()
{
    for (auto& s : spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, s.a, s.b) == cudaSuccess) { ms[s.cls] += t; timed[s.cls] += 1; }
        cudaEventDestroy(s.a);
        cudaEventDestroy(s.b);
    }
    spans.clear();
}

LaunchScope::LaunchScope(int cls, cudaStream_t st_, double flop) : p(g_prof), st(st_)
{
    if (!p) return;
    p->launches[cls] += 1;
    p->flops[cls] += flop;
    if (p->timing) {
        Profiler::Span s{cls, nullptr, nullptr};
        cudaEventCreate(&s.a);
        cudaEventCreate(&s.b);
        cudaEventRecord(s.a, st);
        p->spans.push_back(s);
        idx = (int)p->spans.size() - 1;
    }
}

LaunchScope::~LaunchScope()
{
    if (p && idx >= 0) cudaEventRecord(p->spans[idx].b, st);
}

} // namespace mcb
