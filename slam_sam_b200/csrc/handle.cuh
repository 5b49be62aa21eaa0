// The handle behind ndtb_handle: owns device memory, one stream, the voxel map and evaluation buffers.
#pragma once
#include "../../include/ndtb.h"
#include "build_api.cuh"
#include "eval_api.cuh"
#include <map>
#include <string>
#include <vector>

namespace ndtb { struct MgState; }  // multigrid.cu

struct ndtb_handle_s {
  int device = 0;
  cudaStream_t stream = nullptr;
  ndtb_params params{};
  ndtb_svn_params svn{};
  std::string last_error;

  // clouds (device, packed float4)
  float4* d_target = nullptr; size_t n_target = 0, target_alloc = 0; bool target_owned = true;
  float4* d_source = nullptr; size_t n_source = 0, source_alloc = 0; bool source_owned = true;
  void* h_stage = nullptr; size_t stage_alloc = 0;  // pinned staging for strided host clouds
  unsigned char* d_raw = nullptr; size_t raw_alloc = 0;   // strided host clouds land here as ONE contiguous copy and are repacked on the device
  // ndtb_stage_target / ndtb_stage_source (stage_api.cu): the NEXT clouds are copied to the device on a copy stream while the
  // current align runs; a FIFO of two per kind, committed by swapping buffers with d_target / d_source
  struct StagedCloud { float4* d = nullptr; size_t alloc = 0, n = 0; cudaEvent_t ready = nullptr, reusable = nullptr; unsigned char* d_raw = nullptr; size_t raw_alloc = 0; };
  StagedCloud staged[2][2];             // [0] target, [1] source
  int staged_head[2] = {0, 0}, staged_count[2] = {0, 0};
  cudaStream_t stage_stream = nullptr;

  // keyframe archive (pointsArchive of the reference's pipelines): clouds uploaded once, windows assembled on device
  struct Keyframe { float4* d_points = nullptr; size_t n = 0; };
  std::map<int64_t, Keyframe> keyframes;

  // ndtb_align_batch scratch, kept between calls (allocation costs milliseconds): four scan slots + a copy stream
  struct BatchSlot { float4* d_src = nullptr; size_t alloc = 0; cudaEvent_t uploaded = nullptr; };
  BatchSlot batch_slot[4];
  cudaStream_t batch_copy_stream = nullptr;
  double* batch_h_results = nullptr;    // pinned, 4 x 2 x (RED_N + 2) words

  // pcl::VoxelGrid down-sampling scratch (ndtb_voxelgrid_filter): its own sort buffers, so the current map is untouched
  ndtb::DeviceGrid ds_grid;
  float4* d_ds_in = nullptr; size_t ds_in_alloc = 0; bool ds_in_owned = true;
  float4* d_ds_out = nullptr; size_t ds_out_alloc = 0;

  // multigrid NDT (pclomp::MultiGridNormalDistributionsTransform): one voxel map per target id, see multigrid.cu
  ndtb::MgState* mg = nullptr;

  ndtb::DeviceGrid grid;
  float built_resolution = 0.f; int built_min_pts = 0; int built_flavour = -1; double built_eig = 0;

  ndtb::EvalBuffers eval;
  ndtb::EvalState* h_states = nullptr;  // pinned, [max_particles]
  double* h_results = nullptr;          // pinned, [max_particles][RED_N]; the pclomp kernel writes it directly (zero-copy)
  volatile unsigned int* h_flag = nullptr;  // pinned completion flag the pclomp kernel publishes (host polls it)
  unsigned int flag_seq = 0;
  std::vector<cudaEvent_t> ev_pool;     // per-evaluation event pairs, resolved after the loop (no per-iteration sync)
  size_t ev_used = 0;
  bool kernel_events = false;           // ndtb_set_profiling: CUDA-event pair around every evaluation (one launch each)
  // alignment session (align_session_kernel): mailbox the resident kernel polls, device-side command copy
  ulonglong2* h_mailbox = nullptr;      // pinned, SESSION_UNITS x 16 bytes
  unsigned long long* d_cmd = nullptr;  // device, SESSION_DCMD_WORDS
  unsigned long long mail_seq = 0;      // commands sent so far (never repeats: keys a stale mailbox cannot satisfy)
  bool session_open = false;
  bool use_sessions = true;             // off with per-kernel events (ndtb_set_profiling) or NDTB_NO_SESSION=1
  double session_idle_ms = 50.0;        // the resident kernel leaves after this long without a command
  bool batch_interleave = true;         // ndtb_align_batch keeps three alignments in flight (ndtb_set_option "batch_interleave")
  bool direct_solver = true;            // Newton direction by pivoted elimination when well conditioned (ndtb_set_option)
  // device Newton step inside the session kernel (SESSION_SPECULATE, derivatives.cu): the host repeats every step and compares
  ulonglong2* h_spec = nullptr;         // pinned, SESSION_SPEC_AREA_UNITS x 16 bytes: records, result-slot sets and the host's acknowledgement (eval_api.cuh)
  bool device_newton = false;           // ndtb_set_option "device_newton": opt-in, the host step is faster
  bool session_spec = false;            // the resident kernel that is open was built with the step (align_session_kernel<., true>)
  bool spec_armed = false;              // the command of the evaluation in flight allows the device to go on by itself
  struct SpecCarry { double a_t = 0; int nr_iterations = 0; int first = 0; bool valid = false; } spec_carry;  // set by align_impl before each evaluation
  unsigned long long spec_accepted = 0, spec_declined = 0, spec_mismatches = 0;  // life-time counters (ndtb_get_session_stats)
  int spec_fault_at = 0, spec_checks = 0;   // test hook: the spec_fault_at-th comparison of an align is treated as a mismatch
  unsigned long long seq_init[3] = {0, 0, 0};
  volatile unsigned long long* pending_time_slots = nullptr;  // the result-slot set that evaluation was published to
  unsigned long long pending_time_key = 0;  // result key of the last session evaluation whose duration slot is unread
  double session_eval_ns = 0;           // device-side duration of the session evaluations of the last align

  bool has_reg_pose = false;
  float reg_translation[3] = {0, 0, 0};

  std::vector<double> trace;  // 50 doubles per evaluation of the last align
  double last_hits = 0, last_found = 0;

  // timing
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_b[2] = {nullptr, nullptr};  // around the last voxel-map build
  bool build_time_pending = false, align_time_pending = false;  // elapsed times are resolved by ndtb_get_timings, off the critical path
  float build_ms = 0, align_ms = 0, eval_ms_total = 0; int eval_launches = 0;
  uint64_t launches = 0;  // eval / stein / score kernels launched (builds are counted in grid.launches)

  // ---- SVN-NDT, device-resident loop (svn_stein.cu, svn_api.cu) -------------------------------------------------
  // Particles, records and updates stay in HBM for the whole align; one iteration = state kernel (poses -> EvalState of the
  // rank's own particles) + eval kernel (its last blocks store the records into every rank's window) + Stein kernel (acquires
  // the K flags, updates all K particles, mean pose + stop test in its last block).  The host only polls a pinned status slot.
  struct SvnDev {
    static constexpr int CAPACITY = 1024;       // particles per window half
    // exchange window: flags [2][CAPACITY] u64, then records [2][CAPACITY][43] f64 -- ONE cudaMalloc so that it can be shared by
    // cudaIpc (one process per GPU) or reached through peer access (one process, several handles)
    unsigned char* d_window = nullptr;
    double* d_poses[2] = {nullptr, nullptr};    // [CAPACITY][16], ping-pong (the Stein kernel reads all poses while it writes the new ones)
    double* d_updates = nullptr;                // [CAPACITY][6]
    double* d_mean = nullptr;                   // [2][16] mean pose ping-pong + [16] prior
    int* d_done = nullptr;                      // 0 running, 1 converged, 2 peer time-out
    unsigned int* d_ticket = nullptr;           // last-block-done ticket of the Stein kernel
    unsigned long long* d_wait_ns = nullptr;    // time block 0 of the Stein kernel spent waiting for records (summed over iterations)
    volatile unsigned long long* h_status = nullptr;  // pinned, 64 self-validating slots {word, word ^ key(stamp)}
    int cur = 0;                                // d_poses[cur] holds the current particles
    unsigned long long stamp = 0;               // iterations executed so far on this handle (all ranks advance in lock-step)
    unsigned long long local_seq = 0;           // stamps of the host-staged parity hooks (this handle only)
    // communicator
    int rank = 0, world = 1;
    void* peer_base[8] = {};                    // window base of every rank (peer_base[rank] == d_window)
    bool peer_is_ipc[8] = {};
    // per-align timing (CUDA events, resolved after the loop)
    std::vector<cudaEvent_t> ev;
    float eval_ms = 0, stein_ms = 0, wait_ms = 0;
    int iterations = 0;
  } svnd;
};

namespace ndtb {
int ensure_eval_buffers(ndtb_handle_s* h, int particles, int n_points);
int set_error(ndtb_handle_s* h, int code, const char* what);
int upload_cloud(ndtb_handle_s* h, const float* xyz, size_t n, size_t stride_bytes, float4** d_buf, size_t* alloc, bool* owned);
// n points, xyz at every stride_bytes (12 or >= 16, a multiple of 4) -> packed float4 at d_dst on `stream`.  A 16-byte stride is one
// copy; any other stride is ONE contiguous copy into *d_raw (grown as needed) + a repack kernel: a 2-D DMA with 12- or 16-byte rows
// runs at a fraction of the link rate (pcl::PointXYZI, 32 B: 1.69 ms per million points against 0.6 ms; packed xyz: 11 ms against 0.25 ms).
int copy_cloud_to_device(ndtb_handle_s* h, const float* xyz, size_t n, size_t stride_bytes, float4* d_dst, cudaStream_t stream, unsigned char** d_raw,
                         size_t* raw_alloc);
int rebuild_grid(ndtb_handle_s* h);
void fill_eval_state(ndtb_handle_s* h, EvalState& s, const double* p, const float* T16, double snap, bool svn);
void finish_eval(ndtb_handle_s* h, int n, const double* r, const double* p, double* score, double* g, double* H, double* nvtl);
int compute_hessian_f64(ndtb_handle_s* h, const double* p, const float* T16, double* H);  // line_search.cu
int mg_eval(ndtb_handle_s* h, const double* p, const float* T16, bool compute_hessian, double* score, double* g, double* H, double* nvtl);  // multigrid.cu
bool mg_ready(ndtb_handle_s* h);
void mg_free(ndtb_handle_s* h);
void svn_dev_free(ndtb_handle_s* h);  // svn_stein.cu
void stage_free(ndtb_handle_s* h);    // stage_api.cu
}  // namespace ndtb
