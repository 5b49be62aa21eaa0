// Host-side interface of the derivative evaluation kernels (derivatives.cu).
#pragma once
#include "ndtb_internal.cuh"

namespace ndtb {

struct EvalBuffers {
  EvalState* d_states = nullptr;   // [max_particles]
  double* d_partials = nullptr;    // [max_particles][max_blocks][RED_N]
  double* d_results = nullptr;     // [max_particles][RED_N]
  unsigned int* d_tickets = nullptr;  // [max_particles] last-block-done counters (self-resetting)
  int max_particles = 0, max_blocks = 0;
  int free_sms = 0;   // DIRECT1 / DIRECT7 kernels (one CTA per SM) leave this many SMs to other streams ("session_free_sms")
};

constexpr int EVAL_THREADS = 128;
#ifndef EVAL_MIN_BLOCKS
#define EVAL_MIN_BLOCKS 3
#endif

// SVN record exchange, performed by the LAST block of every particle's evaluation (svn flavour only): the particle's
// 43-double record {score, g[6], H[36] + 1e-6 I, NaN guards of svn_ndt_impl.hpp:650-663} is stored straight into the record
// window of EVERY rank (rec[r] are peer-mapped device pointers over NVLink, rec[rank] is the local window), followed by a
// system-scope release of the particle's flag.  The consumer (stein_kernel) acquires the K flags.  No collective call, no
// host in the loop.  n_dest == 0 disables the exchange (results only go to EvalBuffers::d_results).
constexpr int SVN_MAX_RANKS = 8;
constexpr int SVN_REC = 43;
struct SvnExchange {
  int world = 1, rank = 0;          // block row s evaluates UNIT u = rank + s * world: particle u / vshards, point shard u % vshards
  int vshards = 1;                  // point shards per particle (1: a unit is a whole particle and the record carries the guards of
                                    // svn_ndt_impl.hpp:650-663; > 1: raw partial sums, the consumer adds the shards in order first)
  int blocks_per_unit = 0;          // vshards > 1: CTAs per unit (0 = eval_blocks_for(n) / vshards)
  int n_dest = 0;                   // windows the record is stored to: rec[0..n_dest-1] (n_dest == world across GPUs, 1 = local only)
  int capacity = 0;                 // particles per parity half of a window
  int parity = 0;                   // window half of this iteration (stamp & 1)
  unsigned long long stamp = 0;     // value the particle's flag is set to
  const int* done = nullptr;        // device flag: != 0 -> the align has finished, the launch is a no-op
  double* rec[SVN_MAX_RANKS] = {};              // [2][capacity][43]
  unsigned long long* flag[SVN_MAX_RANKS] = {}; // [2][capacity]
};

// One derivative evaluation for `n_particles` poses over `n` source points.
//   flavour  : FLAVOUR_PCLOMP -> ndt_omp_impl.hpp:217-347 maths (float pair maths, FP64 accumulate)
//              FLAVOUR_SVN    -> svn_ndt_impl.hpp:518-647 maths (FP64 Mahalanobis/exp, GN or full Hessian)
//   n_probe  : 7 (DIRECT7) or 1 (DIRECT1)
// Results land in bufs.d_results[particle][RED_N] on `stream`.
void launch_eval(int flavour, int n_probe, bool compute_hessian, bool svn_gauss_newton, const float4* d_src, int n, const GridMeta& meta,
                 const VoxRec* recs, const HashSlot* table, const double* c_mean, const double* c_icov, const EvalBuffers& bufs,
                 int n_particles, int particle_offset, cudaStream_t stream, const SvnExchange* xch = nullptr);

int eval_blocks_for(int n);

// pclomp flavour, ONE pose passed by value (kernel parameter, no state upload).  The 46 results are
// written to `results` (device or mapped pinned host memory) as self-validating 16-byte slots
// {bits(value), bits(value) ^ result_key}: a host thread can poll the slots instead of synchronising
// the stream, and needs neither a flag nor a system-scope fence in the kernel.
cudaError_t launch_eval_pclomp(int n_probe, bool compute_hessian, const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs,
                               const HashSlot* table, const EvalState& st, const EvalBuffers& bufs, double* results,
                               unsigned long long result_key, cudaStream_t stream, unsigned long long* dbg_stamps = nullptr);
int eval_blocks_pclomp(int n, int free_sms = 0);

// Alignment session (see align_session_kernel): one command per evaluation, written by the host into a pinned
// mailbox of 64 units {word, word ^ session_key(seq)}; seq increases monotonically over the life of a handle.
struct SessionCmd {
  EvalState st;
  unsigned long long result_key;  // key of the result slots of this evaluation
  unsigned long long flags;       // SESSION_QUIT, SESSION_SPECULATE
  // SESSION_SPECULATE: what the last CTA of this evaluation needs to take the Newton step itself and start the next evaluation
  // without waiting for the host (device_newton.cuh); the host repeats the step and compares, see eval_pclomp()
  double step_max, step_min, trans_eps;
  int max_iterations;
  int first;                      // 1: this is the evaluation at the initial guess
  double a_t;                     // step length that led to this evaluation
  unsigned long long nr_iterations;
  unsigned long long key_seq;     // result_key == result_key_of(key_seq); the next evaluation's is result_key_of(key_seq + 1)
};
constexpr int SESSION_CMD_WORDS = static_cast<int>(sizeof(SessionCmd) / 8);
constexpr int SESSION_FLAGS_WORD = static_cast<int>((sizeof(EvalState) + 8) / 8);
constexpr int SESSION_UNITS = 64;           // mailbox units (16 bytes each)
constexpr int SESSION_SEQ_WORD = 64;        // d_cmd[64]: sequence number released by CTA 0 (or by the last CTA of a speculating evaluation)
constexpr int SESSION_TSEEN_WORD = 65;      // d_cmd[65]: globaltimer when CTA 0 saw the command
constexpr int SESSION_DECLINE_WORD = 66;    // d_cmd[66]: sequence number the last CTA will NOT release itself (CTA 0 polls the mailbox for it)
constexpr int SESSION_DCMD_WORDS = 80;
constexpr int SESSION_SPEC_STATUS_UNIT = 62;  // unit of a speculation record that holds the Newton status
// Speculation area (pinned host memory, 16-byte units), everything double-buffered by the parity of the evaluation's key_seq so
// that the device may run at most two evaluations ahead of the host:
//   units [0, 128)    two records: the command the device built for evaluation key_seq + its status
//   units [128, 224)  two result-slot sets (RED_N + 1 self-validating slots each) of the session's evaluations
//   unit  224         the host's acknowledgement (one 8-byte word s): it has read the results of every evaluation up to key_seq s
constexpr int SESSION_SPEC_RESULT_UNIT0 = 128, SESSION_SPEC_RESULT_UNITS = 48, SESSION_SPEC_ACK_UNIT = 224, SESSION_SPEC_AREA_UNITS = 256;
static_assert(SESSION_SPEC_RESULT_UNITS >= RED_N + 1, "a result-slot set holds RED_N values and the duration");
constexpr unsigned long long SESSION_QUIT = 1ull;
constexpr unsigned long long SESSION_SPECULATE = 2ull;
static_assert(sizeof(SessionCmd) % 8 == 0 && SESSION_CMD_WORDS <= SESSION_SPEC_STATUS_UNIT, "command must fit the mailbox");
static_assert(sizeof(EvalState) % 8 == 0, "the flags word is addressed by index");
__host__ __device__ inline unsigned long long result_key_of(unsigned long long key_seq) { return (0x9E3779B97F4A7C15ull * key_seq) | 1ull; }
__host__ __device__ inline unsigned long long session_key(unsigned long long seq) { return (0xD1B54A32D192ED03ull * seq) | 1ull; }

cudaError_t launch_align_session(int n_probe, const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs, const HashSlot* table,
                                 const ulonglong2* mailbox, unsigned long long* d_cmd, const EvalBuffers& bufs, double* results,
                                 ulonglong2* spec_record, unsigned long long seq0, unsigned long long idle_timeout_ns, cudaStream_t stream);

}  // namespace ndtb
