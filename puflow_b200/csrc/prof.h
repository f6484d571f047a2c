// Optional per-launch timing of the one-call entry points (pf_upsample_patches, pf_upsample_clouds): while enabled, every
// launch is followed by a CUDA event on the caller's stream and the interval since the previous event is recorded under
// the launch's name.  The recorder is thread_local: each host thread profiles its own calls, nothing is shared (the
// library stays re-entrant: one host thread per GPU / stream).  Off by default; costs one branch
// per launch when off.  Do not enable during CUDA-graph capture.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

namespace pf {

struct Recorder {
    bool on = false;
    int depth = 0;                       // nesting of entry points (pf_upsample_clouds calls pf_upsample_patches)
    std::vector<cudaEvent_t> ev;         // event pool
    size_t used = 0;
    struct Interval { std::string name; size_t e0, e1; };
    std::vector<Interval> iv;
    size_t last = 0;

    cudaEvent_t take() {
        if (used == ev.size()) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) { on = false; return nullptr; }
            ev.push_back(e);
        }
        return ev[used++];
    }
    void begin(cudaStream_t st) {        // start of an entry point
        if (!on) return;
        if (depth++ == 0) {
            cudaEvent_t e = take();
            if (!e) return;
            last = used - 1;
            cudaEventRecord(e, st);
        }
    }
    void mark(cudaStream_t st, const char* name) {   // after a launch
        if (!on || depth == 0) return;
        cudaEvent_t e = take();
        if (!e) return;
        cudaEventRecord(e, st);
        iv.push_back(Interval{name, last, used - 1});
        last = used - 1;
    }
    void end() { if (on && depth > 0) depth--; }
    void reset() { used = 0, depth = 0, iv.clear(); }
};

Recorder& recorder();                    // this host thread's recorder (model.cu)

struct ProfScope {                       // begin / end around an entry point, marks in between
    Recorder& r;
    cudaStream_t st;
    explicit ProfScope(cudaStream_t s) : r(recorder()), st(s) { r.begin(st); }
    ~ProfScope() { r.end(); }
    void mark(const char* name) { r.mark(st, name); }
};

}  // namespace pf
