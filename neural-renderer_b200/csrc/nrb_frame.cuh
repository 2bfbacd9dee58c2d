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
    int literalBounceTraces;  // NRB_OPT_LITERAL_BOUNCE_TRACES
    int bounceEntryVote;      // NRB_OPT_BOUNCE_ENTRY = 1: the fused three-class walk under the voting scheduler
    int literalSegExits;      // the option above, or an outer shell that is not a closed surface: k_seg_exits traces the
                              // outer shell to the end, so the reference's exitT = 1e-8 fallback for rays that find no
                              // back face (:899-905) is reproduced exactly
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
StageProfiler* get_stage_profiler();
void debug_check(const char* what, cudaStream_t s);
bool debug_sync_enabled();   // NRB_DEBUG in the environment, read once per process
// Multiprocessor count of the CURRENT device (cached per device ordinal): persistent grids are sized from it.
int sm_count();
// Per-device opt-in to > 48 KB of dynamic shared memory for every kernel that needs it.  Function attributes
// belong to the device/context, so nrb_create runs these once for each device it is given (a process may hold
// renderers on several GPUs).
void init_frame_kernels();      // nrb_segments.cu
void init_query_kernels();      // nrb_neural.cu (+ nrb_neural_tc.cu, nrb_query_split.cu)
void init_query_tc_kernels();
void init_query_split_kernels();
#define NRB_LAUNCHED(what, s) ::nrb::debug_check(what, s)

// Launchers implemented in nrb_frame.cu / nrb_neural.cu (all enqueue on `s`, none synchronises).
struct NetDev;
void launch_trace_rays(const MeshDev& m, int mode, const float* o, const float* d, size_t n, Hit* out,
                       unsigned long long* stats, cudaStream_t s);
int launch_gt_frame(const Frame& f, bool lambert, cudaStream_t s, int* hostFeedback, int pathTailBelow,
                    cudaStream_t sOut = nullptr, cudaEvent_t evSplit = nullptr);  // returns kernels launched; sOut: see launch_neural_frame
// Host-side state of the device-driven segment-loop tail (nrb_segments.cu): one instantiated
// conditional-WHILE graph per hit-record target, the device copy of the frame block its kernels
// read, and the key the graphs were built for.  Owned by the renderer; null = fixed ten iterations.
struct SegTail {
    cudaGraph_t graph[2] = {nullptr, nullptr};
    cudaGraphExec_t exec[2] = {nullptr, nullptr};
    cudaStream_t capture = nullptr;
    void* dFrame = nullptr;       // Frame
    void* dCtl = nullptr;         // SegCtl[2]
    unsigned char key[512] = {};  // bytes the graphs depend on (NetDev, state pointers, sizes)
    size_t keyBytes = 0;
    bool failed = false;          // conditional graphs unavailable: stay on the fixed loop
    void release();
    ~SegTail() { release(); }
};
// sOut / evSplit: frame pipelining - with sOut != s the last part of the frame (k_path_tail, finalize) is enqueued on sOut
// behind an event recorded on s; nullptr = everything on s.
int launch_neural_frame(const Frame& f, const NetDev& net, bool lambert, cudaStream_t s, SegTail* tail, bool fusedTail,
                        int* hostFeedback, int pathTailBelow, cudaStream_t sOut = nullptr, cudaEvent_t evSplit = nullptr);
size_t tile_sort_temp_bytes(int n);
int launch_tile_sort(const unsigned int* cost, unsigned int* costSorted, int* ids, int* order, int n, void* temp, size_t tempBytes,
                     cudaStream_t s);
int launch_finalize(const Frame& f, cudaStream_t s);
int launch_lambert(const Frame& f, cudaStream_t s);
int launch_denoise(const Frame& f, cudaStream_t s);
// peer-memory frame gather hand-shake (nrb_frame.cu)
void launch_gather_release(unsigned* sync, unsigned frames, cudaStream_t s);
void launch_gather_wait(const unsigned* flag, unsigned want, unsigned* err, cudaStream_t s);
void launch_gather_signal(unsigned* remoteSync, cudaStream_t s);  // world == 1 only: reads accum/rec[0] of the whole frame

}  // namespace nrb
