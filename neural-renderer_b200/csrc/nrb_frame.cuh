// nrb_frame.cuh — the kernel-argument block shared by all frame kernels and their launchers.
#pragma once

#include "nrb_state.cuh"

namespace nrb {

// Passed by value as a __grid_constant__ kernel parameter (lives in the constant bank; the
// reference passes RenderParams + MeshDeviceView + a dozen pointers the same way).
struct Frame {
    FrameParams fp;
    MeshDev mesh[4];
    EnvDev env;
    MaterialDev gmat;
    State st;
    int classicSlot;
    int additionalEnabled;
};

// Per-launch hook.  NRB_DEBUG=1 in the environment: synchronise + report after every launch.
// With a StageProfiler installed (NRB_OPT_PROFILE): record a CUDA event after every launch so the
// host can attribute device time to kernels (the reference's PROF_BEGIN/PROF_END pairs,
// src/cuda_renderer_neural.cu:39-57, always compiled in here).
struct StageProfiler {
    static constexpr int kMaxMarks = 1024;
    cudaEvent_t ev[kMaxMarks + 1];
    const char* name[kMaxMarks];
    int marks = 0;
    bool ready = false;
};
void set_stage_profiler(StageProfiler* p);
void debug_check(const char* what, cudaStream_t s);
#define NRB_LAUNCHED(what, s) ::nrb::debug_check(what, s)

// Launchers implemented in nrb_frame.cu / nrb_neural.cu (all enqueue on `s`, none synchronises).
struct NetDev;
void launch_trace_rays(const MeshDev& m, int mode, const float* o, const float* d, size_t n, Hit* out,
                       unsigned long long* stats, cudaStream_t s);
int launch_gt_frame(const Frame& f, bool lambert, cudaStream_t s);       // returns kernels launched
int launch_neural_frame(const Frame& f, const NetDev& net, bool lambert, cudaStream_t s);
int launch_finalize(const Frame& f, cudaStream_t s);
int launch_lambert(const Frame& f, cudaStream_t s);

}  // namespace nrb
