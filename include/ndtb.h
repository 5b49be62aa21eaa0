/* ndtb.h -- C ABI of the B200-native NDT scan-registration engine (libndtb.so).
 *
 * This is the drop-in boundary for ONE hot path of khalisfadil/slam-sam: NDT scan registration
 * (target voxel-map build, per-point score/gradient/Hessian, Newton loop, SVN-NDT).  The reference
 * has no FFI layer: the path sits behind two C++ header-template classes.  Each entry point below
 * cites the reference method(s) it replaces; include/ndtb_pclomp.hpp re-creates the reference's
 * class/method names on top of this ABI (see INTEGRATION.md for the binding a maintainer adds).
 *
 * Conventions: every function returns an int status (NDTB_OK == 0, negative = error); no C++
 * exception crosses the ABI; the caller owns all host buffers; a handle owns its device memory and
 * one CUDA stream; one handle per caller thread (like the reference's registrars, which are not
 * re-entrant).  4x4 matrices are ROW-MAJOR (m[4*r+c]); 6-vectors are (x,y,z,roll,pitch,yaw) unless
 * stated; 6x6 matrices row-major.  There is NO CPU fallback: without a CUDA device every compute
 * entry point fails with NDTB_ERR_CUDA.
 */
#ifndef NDTB_H_
#define NDTB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NDTB_OK 0
#define NDTB_ERR_INVALID (-1)   /* bad argument / handle                                           */
#define NDTB_ERR_CUDA (-2)      /* CUDA runtime error (see ndtb_last_error)                         */
#define NDTB_ERR_EMPTY (-3)     /* empty / missing cloud (reference: PCL_WARN/PCL_ERROR + early out)*/
#define NDTB_ERR_OVERFLOW (-4)  /* dx*dy*dz > INT32_MAX: voxel_grid_covariance_omp_impl.hpp:83-94   */
#define NDTB_ERR_NOT_READY (-5) /* target or source not set                                         */

/* pclomp::NeighborSearchMethod values (extern/ndt_omp/include/pclomp/ndt_struct.hpp) */
#define NDTB_KDTREE 0   /* radius search over the leaf means (27 hash probes + FLANN's float distance test); pclomp flavour: derivatives + align without line search; svn flavour: svn_ndt_impl.hpp:576-579 */
#define NDTB_DIRECT26 1 /* pclomp flavour, derivatives + align without line search (generic kernel, off the hot path)      */
#define NDTB_DIRECT7 2
#define NDTB_DIRECT1 3

#define NDTB_FLAVOUR_PCLOMP 0 /* pclomp::VoxelGridCovariance + pclomp NDT maths   */
#define NDTB_FLAVOUR_SVN 1    /* svn_ndt::VoxelGridCovariance + svn_ndt maths     */

typedef struct ndtb_handle_s* ndtb_handle;

/* Mirrors pclomp::NdtParams (ndt_struct.hpp:90-105) + the knobs that live elsewhere in the
 * reference: outlier_ratio_ (ndt_omp_impl.hpp:54), min_points_per_voxel_ and
 * min_covar_eigvalue_mult_ (voxel_grid_covariance_omp.h:193-208), PointCloud::is_dense. */
typedef struct ndtb_params {
  float resolution;                  /* setResolution            ndt_omp.h:135-142            */
  double step_size;                  /* setStepSize              ndt_omp.h:159                */
  double trans_epsilon;              /* setTransformationEpsilon (pcl::Registration)          */
  int max_iterations;                /* setMaximumIterations     (pcl::Registration)          */
  int search_method;                 /* setNeighborhoodSearchMethod ndt_omp.h:121             */
  double outlier_ratio;              /* setOulierRatio           ndt_omp.h:173                */
  int min_points_per_voxel;          /* VoxelGridCovariance::setMinPointPerVoxel              */
  double min_covar_eigvalue_mult;    /* VoxelGridCovariance::setCovEigValueInflationRatio     */
  float regularization_scale_factor; /* setRegularizationScaleFactor ndt_omp.h:275            */
  int use_line_search;               /* NdtParams::use_line_search (More-Thuente), default 0  */
  int grid_flavour;                  /* NDTB_FLAVOUR_*                                        */
  int is_dense;                      /* target cloud is_dense (non-finite points skipped if 0)*/
} ndtb_params;

/* pclomp::NdtResult (ndt_struct.hpp:67-88) + hasConverged(). */
typedef struct ndtb_result {
  float pose[16];                 /* getFinalTransformation()                                    */
  double hessian[36];             /* getHessian()                                                */
  double transform_probability;   /* getTransformationProbability()                              */
  double nearest_voxel_transformation_likelihood;
  int iteration_num;              /* getFinalNumIteration()                                      */
  int converged;                  /* hasConverged()                                              */
  int transformation_array_len;   /* entries written to `transformation_array` (<= capacity)     */
  int evaluations;                /* computeDerivatives calls made (= roofline "E")              */
} ndtb_result;

/* svn_ndt::SvnNdtResult (extern/svn_ndt/include/svn_ndt.h:40-51). */
typedef struct ndtb_svn_params {
  int particle_count;        /* setParticleCount        svn_ndt.h:112 */
  int max_iterations;        /* setMaxIterations                       */
  double kernel_bandwidth;   /* setKernelBandwidth                     */
  double step_size;          /* setStepSize                            */
  double stop_threshold;     /* setEarlyStopThreshold                  */
  int use_gauss_newton;      /* setUseGaussNewtonHessian               */
} ndtb_svn_params;

typedef struct ndtb_svn_result {
  double final_pose[16];        /* gtsam::Pose3 as 4x4                                        */
  double final_covariance[36];  /* GTSAM tangent order (rx,ry,rz,x,y,z)                       */
  int converged;
  int iterations;
} ndtb_svn_result;

/* ---- lifetime ------------------------------------------------------------------------------- */
int ndtb_create(int cuda_device, ndtb_handle* out); /* ctor: ndt_omp_impl.hpp:51-81 defaults     */
int ndtb_destroy(ndtb_handle h);
const char* ndtb_last_error(ndtb_handle h);
const char* ndtb_version(void);

/* ---- parameters ----------------------------------------------------------------------------- */
int ndtb_default_params(int grid_flavour, ndtb_params* p); /* ctor defaults of either flavour    */
int ndtb_set_params(ndtb_handle h, const ndtb_params* p);  /* rebuilds the grid iff resolution /
                                                              min points / flavour changed and a
                                                              target is set (ndt_omp.h:135-142)  */
int ndtb_get_params(ndtb_handle h, ndtb_params* p);

/* ---- clouds (host pointers; xyz = 3 consecutive floats at each `stride_bytes`) --------------- */
/* setInputTarget: ndt_omp.h:126-130 -> init() :324-330 -> VoxelGridCovariance::filter(true)
 * (voxel_grid_covariance_omp_impl.hpp:55-364) or svn_ndt_impl.hpp:137-159. Builds the voxel map. */
int ndtb_set_target(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes);
/* setInputSource (pcl::Registration). */
int ndtb_set_source(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes);
/* Same, but the buffers already live in device memory of the handle's GPU (packed float4). */
int ndtb_set_target_device(ndtb_handle h, const void* d_xyzw, size_t n);
int ndtb_set_source_device(ndtb_handle h, const void* d_xyzw, size_t n);
/* Pipelined form of setInputTarget / setInputSource (ndt_omp.h:126-130; the reference's clouds are host memory, so it has no
 * copy to hide): ndtb_stage_* starts the host-to-device copy of the NEXT cloud on a copy stream and returns at once (the host
 * buffer must stay valid, and should be pinned, until the matching commit); ndtb_commit_* makes the oldest staged cloud the
 * current one -- ndtb_commit_target then builds the voxel map exactly like ndtb_set_target.  Up to two clouds of each kind may
 * be staged (NDTB_ERR_NOT_READY beyond that, or when committing with nothing staged), so cloud i + 1 can be staged before
 * cloud i is committed and its copy overlaps the build and align of cloud i. */
int ndtb_stage_target(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes);
int ndtb_stage_source(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes);
int ndtb_commit_target(ndtb_handle h);
int ndtb_commit_source(ndtb_handle h);
int ndtb_staged_count(ndtb_handle h, int* targets, int* sources);
int ndtb_stage_sync(ndtb_handle h); /* blocks until every staged copy has arrived (a commit does not need it) */
/* setRegularizationPose / unsetRegularizationPose: ndt_omp.h:277-283.  T == NULL unsets. */
int ndtb_set_regularization_pose(ndtb_handle h, const float* T16);

/* ---- pclomp NDT ------------------------------------------------------------------------------ */
/* computeDerivatives (ndt_omp_impl.hpp:217-392) at pose p, cloud transformed by T(p) like
 * computeStepLengthMT :969-983.  If T16 != NULL that float matrix transforms the cloud instead
 * (the initial-guess evaluation :101-135).  Parity hook for SURVEY 8(a) row a8. */
int ndtb_compute_derivatives(ndtb_handle h, const double p[6], const float* T16, int compute_hessian, double* score,
                             double gradient[6], double hessian[36], double* nvtl);
/* ndtb_compute_derivatives without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernel and same results as ndtb_compute_derivatives (DIRECT1 / DIRECT7, pclomp flavour), but the
 * kernel's scratch -- per-block partial sums, last-block ticket, result slots, 55 KB -- is a static __device__ array
 * inside the library (every GPU that loads the library holds it until exit).  The call performs no cudaMalloc /
 * cudaFree / allocating library call: the target map must already be built (ndtb_set_target*), the source is
 * `d_source_xyzw` (device memory, n_source packed float4 points, w unused) or, when NULL, the handle's resident source.
 * n_source > NDTB_STATIC_MAX_POINTS is rejected with NDTB_ERR_INVALID.  Outputs are host pointers as in
 * ndtb_compute_derivatives.  All calls on one GPU share the scratch: the call clears what must start at zero, returns
 * after the result is back on the host, and calls that overlap on the same GPU (other threads / handles) race. */
#define NDTB_STATIC_MAX_POINTS 4194304
int ndtb_compute_derivatives_static(ndtb_handle h, const void* d_source_xyzw, size_t n_source, const double p[6], const float* T16,
                                    int compute_hessian, double* score, double gradient[6], double hessian[36], double* nvtl);
/* align(output, guess): pcl::Registration::align -> computeTransformation (ndt_omp_impl.hpp:85-202).
 * `out_xyz` (optional, n_source*3 floats) receives the transformed source like `output`.
 * `transformation_array` (optional, capacity*16 floats) receives transformation_array_. */
int ndtb_align(ndtb_handle h, const float guess[16], ndtb_result* result, float* transformation_array, int capacity, float* out_xyz);
/* Per-evaluation trace of the last ndtb_align: 50 doubles per evaluation (p6, score, g6, H36, nvtl). */
int ndtb_get_trace(ndtb_handle h, double* trace, int capacity, int* count);
/* calculateTransformationProbability / calculateNearestVoxelTransformationLikelihood
 * (ndt_omp_impl.hpp:1169-1281): FP64 score-only pass over a cloud transformed by T16. */
int ndtb_calculate_scores(ndtb_handle h, const float T16[16], double* transformation_probability, double* nvtl);

/* calculateScore (ndt_omp_impl.hpp:1083-1167), from cleared maps (cleanScores, ndt_omp.h:236-240): *score = sum / N;
 * getScoreMap() -> voxel_idx[] / voxel_mean_score[] (ascending key = getLeafIndex(cell mean), value = the voxel's mean score over the
 * pairs that hit it); getEmptyVoxels() -> empty_ijk[3 * i] = floor(x / resolution) of the points that found no neighbour, in the
 * set's EigenCmp order (ndt_omp.h:57-65).  *n_voxels / *n_empty report the full counts; at most `capacity` entries are written.
 * nomap_points_num_ is keyed by an uninitialised variable in the reference (:1124-1127) and is not provided. */
int ndtb_calculate_score(ndtb_handle h, const float T16[16], double* score, size_t map_capacity, int64_t* voxel_idx, double* voxel_mean_score,
                         size_t* n_voxels, size_t empty_capacity, int32_t* empty_ijk, size_t* n_empty);

/* ---- voxel map inspection (getTargetCells().getLeaves(), include/pipeline.hpp:163-222) -------- */
int ndtb_get_voxel_count(ndtb_handle h, size_t* count, int min_b[3], int max_b[3], int div_b[3]);
/* Arrays sized by count, ascending linear index: idx[V] nr_points[V] mean[3V] cov[9V] icov[9V]
 * evecs[9V] evals[3V]; any pointer may be NULL. */
int ndtb_get_voxels(ndtb_handle h, int64_t* idx, int32_t* nr_points, double* mean, double* cov, double* icov, double* evecs,
                    double* evals);
/* extractNdtData + writeNdtDataToFiles (include/pipeline.hpp:163-274), the NDT map export of the reference's pipelines:
 * `ellipsoid_filename` (ndt_ellipsoids.txt) one line per exported leaf: mean(3) evecs(9, row-major) evals(3) point_count;
 * `voxel_filename` (ndt_voxels.txt): getLeafCenter(index)(3) resolution; std::fixed, 8 decimals, ascending leaf index.  A leaf is
 * exported when `getPointCount() >= min_points` in the reference's int-vs-size_t comparison (:194), so leaves flagged -1 are written
 * too, with the count static_cast<size_t>(-1).  `points_filename` (optional, may be NULL): the caller's map cloud as
 * "x y z intensity" (intensity = fifth float of a >= 20-byte record, else 0).  Host code over the exported voxel table. */
int ndtb_write_ndt_files(ndtb_handle h, const char* ellipsoid_filename, const char* voxel_filename, const char* points_filename,
                         const float* points, size_t n_points, size_t stride_bytes);
/* Mean hits per source point of the last derivative evaluation (roofline h-bar). */
int ndtb_get_last_hits(ndtb_handle h, double* mean_hits_per_point, double* points_with_neighbours);

/* ---- SVN-NDT (grid_flavour must be NDTB_FLAVOUR_SVN) ------------------------------------------ */
int ndtb_svn_default_params(ndtb_svn_params* p);      /* svn_ndt_impl.hpp:61-77                  */
int ndtb_svn_set_params(ndtb_handle h, const ndtb_svn_params* p);
/* computeParticleDerivatives for K poses (svn_ndt_impl.hpp:518-668 under :758-781). poses: K x 16
 * doubles.  Outputs (NDT order): scores[K], grads[6K], hess[36K]. */
int ndtb_svn_particle_derivatives(ndtb_handle h, int K, const double* poses, double* scores, double* grads, double* hess);
/* Stage 2 (svn_ndt_impl.hpp:789-839): updates[6K] in GTSAM order. */
int ndtb_svn_stage2(ndtb_handle h, int K, const double* poses, const double* grads, const double* hess, double* updates);
/* One full iteration (stages 1-3) on K particles, in/out K x 16 doubles. Optional outputs. */
int ndtb_svn_step(ndtb_handle h, int K, double* poses, double* scores, double* grads, double* hess, double* updates);
/* align(source, prior) (svn_ndt_impl.hpp:675-964).  init_particles: K x 16 doubles or NULL (then
 * drawn from `seed` with the reference's sigmas, :710 -- the reference seeds from the clock).
 * particles_out: optional K x 16. */
/* The whole loop runs on the device: per iteration a state kernel, the derivative kernel and one Stein kernel (K x K update,
 * retract, mean pose, stop test) are queued; the host polls one pinned status word per iteration, one iteration behind. */
int ndtb_svn_align(ndtb_handle h, const double prior[16], const double* init_particles, uint64_t seed, ndtb_svn_result* result,
                   double* particles_out);
/* Device time of the last ndtb_svn_align, summed over its iterations: derivative kernels (CUDA events), Stein kernels (events; with
 * a communicator this includes waiting for the peers' records) and the part of it block 0 of the Stein kernel spent waiting for
 * records (in-kernel timer). */
int ndtb_svn_get_timings(ndtb_handle h, float* eval_ms, float* stein_ms, float* exchange_wait_ms, int* iterations);

/* ---- multi-GPU SVN: particles sharded over GPUs, the reference's stage-1 / stage-2 split (svn_ndt_impl.hpp:758-781 independent
 * per particle, :789-839 sums over all particles) --------------------------------------------------------------------------
 * Every rank holds the same target map and source cloud and evaluates the particles k with k % world == rank.  There is no
 * collective library call on the critical path: the LAST BLOCK of a particle's derivative evaluation stores the particle's
 * 43-double record {score, g, H} into the record window of EVERY rank over NVLink (peer-mapped device memory) and releases a
 * per-particle flag; each rank's Stein kernel acquires the K flags and updates all K particles redundantly, so poses never
 * travel and all ranks stay bit-identical.  ~34 KB cross NVLink per iteration at K = 100.
 *   one process per GPU (torchrun, MPI ...): ndtb_comm_window_handle on every rank -> exchange the opaque handles out of band
 *       (any all-gather of *handle_bytes bytes) -> ndtb_comm_connect on every rank -> BARRIER -> aligns
 *   one process, one handle per GPU: ndtb_comm_init(handles, n), then ndtb_svn_align from n threads, one per handle
 * With a communicator, ndtb_svn_align (= ndtb_svn_align_sharded, which insists on one) is COLLECTIVE: all ranks call it with
 * the same prior / particles / parameters.  A rank whose peers do not deliver within 4 s returns NDTB_ERR_CUDA instead of
 * hanging.  ndtb_comm_close returns the handle to single-GPU operation. */
int ndtb_comm_window_handle(ndtb_handle h, void* ipc_handle, size_t capacity_bytes, size_t* handle_bytes);
int ndtb_comm_connect(ndtb_handle h, int rank, int world, const void* all_ipc_handles /* world x handle_bytes, rank order */);
int ndtb_comm_init(ndtb_handle* handles, int n);
int ndtb_comm_close(ndtb_handle h);
int ndtb_comm_info(ndtb_handle h, int* rank, int* world);
int ndtb_svn_align_sharded(ndtb_handle h, const double prior[16], const double* init_particles, uint64_t seed, ndtb_svn_result* result,
                           double* particles_out);

/* Host-staged halves of one iteration (parity hooks, and the building blocks of a caller that wants to move the records itself):
 * this handle evaluates particles k with k % world == rank in stage 1; the caller gathers the (score,g,H) records (43 doubles
 * per particle) between ndtb_svn_stage1_local and ndtb_svn_finish_step. */
int ndtb_svn_stage1_local(ndtb_handle h, int K, const double* poses, int rank, int world, double* records43);
int ndtb_svn_finish_step(ndtb_handle h, int K, double* poses, const double* records43, double* updates);

/* Host helpers of the sharded SVN loop (every rank runs them redundantly on identical particles):
 * mean pose = prior.retract(mean_k Logmap(prior^-1 P_k)) (svn_ndt_impl.hpp:867-878) and the stop
 * test || Logmap(a^-1 b) || (:879-884). */
int ndtb_svn_mean_pose(int K, const double* poses, const double prior[16], double mean16[16]);
int ndtb_svn_pose_distance(const double a[16], const double b[16], double* norm);

/* ---- sliding-window target, assembled on the device ------------------------------------------------------
 * The reference's pipelines keep every keyframe cloud in `pointsArchive` and, per frame, build the NDT target from
 * the last few of them: `pcl::transformPointCloud(*archive.points, *cur, Tb2m.matrix()); *target += *cur;`
 * (run/pipeline_ligo_tc.cpp:519-526, run/pipeline_lo_svn.cpp) followed by setInputTarget.  Here a keyframe cloud is
 * uploaded once (ndtb_keyframe_put), and ndtb_set_target_window transforms the listed keyframes with their CURRENT
 * pose estimates (row-major double 4x4, the arithmetic of pcl::transformPointCloud<PointT,double>), concatenates
 * them in the given order and rebuilds the voxel map -- equivalent to the loop above + setInputTarget, with no
 * per-frame host->device copy of the map window. */
int ndtb_keyframe_put(ndtb_handle h, int64_t id, const float* xyz, size_t n, size_t stride_bytes);
int ndtb_keyframe_drop(ndtb_handle h, int64_t id);
int ndtb_keyframe_count(ndtb_handle h, size_t* count, size_t* total_points);
int ndtb_set_target_window(ndtb_handle h, int n_keyframes, const int64_t* ids, const double* poses16);
/* pcl::VoxelGrid<PointT>::filter with a cubic leaf (apps/align.cpp:65-74, run/pipeline_ins_map_distribution.cpp:324-341,
 * PCL >= 1.14 filters/impl/voxel_grid.hpp): one centroid (float32 sum / count, CentroidPoint) per occupied leaf, in ascending
 * leaf index.  dest 0: centroids are copied to `out` (packed x,y,z,1; at most `capacity` points); dest 1: they become the
 * source cloud; dest 2: they become the target cloud and the voxel map is rebuilt.  *n_out = points produced.  xyz only:
 * intensity and other fields are not carried (the registration never reads them). */
int ndtb_voxelgrid_filter(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes, float leaf, int dest, float* out, size_t capacity,
                          size_t* n_out);
/* The current target cloud as packed x,y,z,w floats (device -> host copy; parity tests of the window assembly). */
int ndtb_get_target_points(ndtb_handle h, float* xyzw, size_t capacity_points, size_t* n_points);

/* ---- multigrid NDT: pclomp::MultiGridNormalDistributionsTransform --------------------------------------------
 * TIER IV's dynamic-map variant of ndt_omp (extern/ndt_omp/include/multigrid_pclomp/): the target is a SET of voxel
 * grids keyed by a string id that come and go as map tiles stream in, the neighbour search is a radius search
 * (radius = resolution) over the leaf centroids of all grids, and score / gradient / Hessian are FP64 end to end.
 * The handle's ndtb_params apply (resolution, step_size, trans_epsilon, max_iterations, outlier_ratio,
 * min_points_per_voxel, min_covar_eigvalue_mult, regularization_scale_factor, use_line_search, is_dense;
 * search_method is ignored: KDTREE only, multigrid_ndt_omp_impl.hpp:197).  The source cloud and the regularisation
 * pose are the handle's (ndtb_set_source, ndtb_set_regularization_pose); the pclomp target of ndtb_set_target is a
 * separate object and is not touched.
 *   ndtb_mg_add_target      addTarget(cloud, id)        multigrid_ndt_omp.h:150-155 (an existing id is replaced)
 *   ndtb_mg_remove_target   removeTarget(id)            multigrid_ndt_omp.h:157 (unknown ids are ignored)
 *   ndtb_mg_create_kdtree   createVoxelKdtree()         multigrid_ndt_omp.h:159-163
 *   ndtb_mg_set_input_target setInputTarget(cloud)      multigrid_ndt_omp.h:143-148 (= addTarget("default") + createVoxelKdtree)
 *   ndtb_mg_get_grid_ids    getCurrentMapIDs()          multigrid_ndt_omp.h:325 (ids joined by '\n')
 *   ndtb_mg_compute_derivatives  computeDerivatives     multigrid_ndt_omp_impl.hpp:338-494
 *   ndtb_mg_align           align() / computeTransformation  multigrid_ndt_omp_impl.hpp:210-323; ndtb_get_trace works as for ndtb_align
 * Evaluating or aligning after add/remove without ndtb_mg_create_kdtree returns NDTB_ERR_NOT_READY (the reference
 * would search a stale kd-tree whose leaf pointers may dangle). */
int ndtb_mg_add_target(ndtb_handle h, const char* grid_id, const float* xyz, size_t n, size_t stride_bytes);
int ndtb_mg_remove_target(ndtb_handle h, const char* grid_id);
int ndtb_mg_create_kdtree(ndtb_handle h);
int ndtb_mg_set_input_target(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes);
int ndtb_mg_get_info(ndtb_handle h, int* n_grids_total, int* n_grids_in_tree, size_t* leaves_in_tree, int* tree_is_current);
int ndtb_mg_get_grid_ids(ndtb_handle h, char* buf, size_t capacity, size_t* needed);
int ndtb_mg_get_leaves(ndtb_handle h, size_t capacity, int32_t* nr_points, double* mean, double* icov, float* centroid_xyz, size_t* count);
int ndtb_mg_compute_derivatives(ndtb_handle h, const double p[6], const float* T16, int compute_hessian, double* score, double gradient[6],
                                double hessian[36], double* nvtl);
/* calculateTransformationProbability / calculateNearestVoxelTransformationLikelihood, multigrid_ndt_omp_impl.hpp:1054-1145 */
int ndtb_mg_calculate_scores(ndtb_handle h, const float T16[16], double* transform_probability, double* nvtl);
int ndtb_mg_align(ndtb_handle h, const float guess[16], ndtb_result* res, float* transformation_array, int capacity, float* out_xyz);

/* ---- scan producer: the reference's LidarCallback on the device ------------------------
 * The step BEFORE setInputSource: Ouster RNG19_RFL8_SIG16_NIR16 lidar packets -> one LidarFrame per revolution
 * (src/lidarcallback.cpp; include/dataframe.hpp LidarFrame), so that scans are born in HBM and the raw packets
 * (24 832 bytes per 16 columns x 128 channels) are the only thing that crosses PCIe.
 *   ndtb_lidar_create        LidarCallback::LidarCallback + Initialize   src/lidarcallback.cpp:17-60, 191-341
 *                            (channel stride not in {1,2,4,8,16} or > pixels_per_column -> NDTB_ERR_INVALID where the
 *                            reference throws std::runtime_error, :41-49)
 *   ndtb_lidar_push_packets  LidarCallback::DecodePacket / DecodePacketRng19   :344-351, 631-878, once per packet in
 *                            order: wrong size or packet_type != 1 -> packet ignored (:633-646); a packet whose frame id
 *                            differs from the running one CLOSES the running frame (:658-669) -- its packets, already in
 *                            HBM, are decoded by one kernel (range / z / vehicle-box / reflectivity filters, channel
 *                            stride, order-preserving compaction in packet, column, channel order).
 *                            LEGACY profile (udp_profile_legacy = 1): DecodePacketLegacy :376-629 -- no packet header, the
 *                            frame id travels in every column header, so a frame is cut at the COLUMN whose id differs (the
 *                            host reads the 12 header bytes of every column; a packet that straddles two frames is copied
 *                            into both), columns need block status 0xFFFFFFFF, range = 20 bits.  When one packet completes
 *                            two frames the reference's `completed_frame` keeps only the later one: so does the pop.  Not
 *                            reproduced: `interframe_timedelta` when one packet holds more than one frame switch, and the
 *                            `frame_id_` reset of :416-418 (it needs a column time stamp of exactly 0.0 s).
 *   ndtb_lidar_pop_frame     the unique_ptr<LidarFrame> DecodePacket returns: oldest closed frame with >= 1 point
 *                            (a frame without points is never emitted, :659); the previously popped frame goes back
 *                            to the pool (ReturnFrameToPool, :903-909).  *available = 0 when there is none.
 *   ndtb_lidar_read_frame    copies the popped frame's arrays to the host (any pointer may be NULL)
 *   ndtb_lidar_frame_points  device pointer to the popped frame's packed (x, y, z, 1) float4 points: pass it to
 *                            ndtb_set_source_device / ndtb_set_target_device (same device); valid until the next pop.
 * Packets handed to ndtb_lidar_push_packets from PINNED host memory are DMA-ed straight from the caller's buffer:
 * keep it unchanged until the frame was popped (or ndtb_lidar_sync returned); pageable buffers may be reused at once. */
typedef struct ndtb_lidar_s* ndtb_lidar;
typedef struct ndtb_lidar_config {
  int columns_per_frame;                /* lidar_data_format.columns_per_frame                       */
  int pixels_per_column;                /* lidar_data_format.pixels_per_column (<= 128 x stride)      */
  int columns_per_packet;               /* config_params.columns_per_packet (<= 32)                   */
  int channel_stride;                   /* lidar_parameter.channelStride                              */
  int reflectivity_threshold;           /* lidar_parameter.reflectionThreshold (stored as uint8_t)    */
  float lidar_origin_to_beam_origin_mm; /* beam_intrinsics.lidar_origin_to_beam_origin_mm             */
  float range_min, range_max;           /* lidar_parameter.rangeFilter                                */
  float z_min, z_max;                   /* lidar_parameter.zAxisFilter                                */
  float box_center[3];                  /* lidar_parameter.vehicleFilterBox.center                    */
  float box_dimensions[3];              /* lidar_parameter.vehicleFilterBox.dimensions                */
  double body_to_lidar_rotation[9];     /* lidar_parameter.Cb2s, row-major                            */
  double body_to_lidar_translation[3];  /* lidar_parameter.tb2s                                       */
  int udp_profile_legacy;               /* config_params.udp_profile_lidar: 0 = "RNG19_RFL8_SIG16_NIR16", 1 = "LEGACY" */
} ndtb_lidar_config;
typedef struct ndtb_lidar_frame_info {  /* scalar members of LidarFrame, include/dataframe.hpp */
  double timestamp;                     /* column time stamp of the first point                       */
  double timestamp_end;                 /* latest column time stamp seen when the frame closed        */
  double interframe_timedelta;
  uint32_t numberpoints;
  uint16_t frame_id;
  uint16_t packets;                     /* packets decoded into this frame (not in the reference)      */
} ndtb_lidar_frame_info;
int ndtb_lidar_default_config(ndtb_lidar_config* cfg); /* member initialisers of include/lidarcallback.hpp:26-90 */
int ndtb_lidar_create(int cuda_device, const ndtb_lidar_config* cfg, const float* beam_azimuth_deg, const float* beam_altitude_deg, ndtb_lidar* out);
int ndtb_lidar_destroy(ndtb_lidar l);
const char* ndtb_lidar_last_error(ndtb_lidar l);
int ndtb_lidar_packet_bytes(ndtb_lidar l, size_t* expected_size); /* expected_size_, :207 */
/* x_1_/y_1_/z_1_ [columns_per_frame x pixels_per_column] and x_2_/y_2_/z_2_ [columns_per_frame] (:283-323) */
int ndtb_lidar_get_tables(ndtb_lidar l, float* x1, float* y1, float* z1, float* x2, float* y2, float* z2);
int ndtb_lidar_push_packets(ndtb_lidar l, const uint8_t* packets, size_t n_packets, size_t packet_bytes, size_t pitch_bytes, int* frames_closed);
int ndtb_lidar_pop_frame(ndtb_lidar l, ndtb_lidar_frame_info* info, int* available);
int ndtb_lidar_read_frame(ndtb_lidar l, size_t capacity, float* x, float* y, float* z, uint16_t* c_id, uint16_t* m_id, double* timestamp_points,
                          float* relative_timestamp, uint16_t* reflectivity, uint16_t* signal, uint16_t* nir);
int ndtb_lidar_frame_points(ndtb_lidar l, const void** d_xyzw, size_t* n_points);
int ndtb_lidar_sync(ndtb_lidar l);
/* kernels launched / CUDA-event time (us) of the last decode kernel */
int ndtb_lidar_get_stats(ndtb_lidar l, uint64_t* launches, float* last_decode_us);

/* ---- batch replay: n independent alignments against the current target ----------------------- */
int ndtb_align_batch(ndtb_handle h, int n_scans, const float* const* xyz, const size_t* n_points, size_t stride_bytes,
                     const float* guesses16, ndtb_result* results);

/* ---- measurement helpers ---------------------------------------------------------------------- */
/* CUDA-event time (ms) of the stages of the last set_target / align on the handle's stream. */
int ndtb_get_timings(ndtb_handle h, float* build_ms, float* align_ms, float* eval_kernel_ms_total, int* eval_kernel_launches);
int ndtb_get_stream(ndtb_handle h, void** cuda_stream);
/* Per-kernel CUDA events around every derivative evaluation: eval_kernel_ms_total of ndtb_get_timings is
 * then the sum of event-pair times, and every evaluation is its own kernel launch.  Off by default: ndtb_align
 * serves all evaluations of a Newton loop from ONE resident kernel (an "alignment session", commands through a
 * pinned mailbox) and eval_kernel_ms_total is the device-side (globaltimer) duration of those evaluations. */
int ndtb_set_profiling(ndtb_handle h, int per_kernel_events);
/* Named switches (no counterpart in the reference):
 *   "direct_solver" (default 1)  Newton direction H^-1(-g) (ndt_omp_impl.hpp:142-145) by elimination with complete
 *                                pivoting + one refinement step when the pivots certify cond(H) <~ 1e6, else -- and
 *                                always with 0 -- by the two-sided Jacobi SVD that restates Eigen::JacobiSVD::solve.
 *                                For a full-rank H the two agree to ~1e-12 relative.
 *   "sessions" (default 1)       0: one kernel launch per evaluation instead of a resident kernel per ndtb_align.
 *   "session_idle_timeout_ms" (default 50)  the resident kernel leaves by itself after this long without a command;
 *                                the host then falls back to one launch per evaluation.
 *   "session_free_sms" (default 0)  the DIRECT1 / DIRECT7 evaluation kernels -- one CTA per SM, the resident kernel of an
 *                                align included -- leave this many of the 148 SMs to other streams, so that the lidar decode
 *                                or the window assembly of the NEXT frame can run while an align is open.  The per-CTA
 *                                partial sums change with the CTA count: results differ from the default by FP64 rounding.
 *   "device_newton" (default 0)  1: the resident kernel of an align takes the Newton step between two evaluations itself (direction
 *                                by pivoted elimination, step clamp, pose -> T, angle tables: ndt_omp_impl.hpp:128-190, 396-538) and
 *                                starts the next evaluation without the PCIe round trip through the host.  The host still
 *                                receives every result, repeats the step with the code the reference path uses (glibc sin / cos)
 *                                and compares the device's next state with its own, word for word; on a difference it lets the
 *                                kernel run out and evaluates its own state with one launch per evaluation, so results do not
 *                                depend on this switch.  Steps that need the host (ill-conditioned H -> SVD, zero slope, the end
 *                                of the loop, a regularisation pose, the line search) are taken by the host as with 0.
 *                                Off by default: measured on a B200 the step costs ~14 us on the device (serial FP64 chains:
 *                                18 dependent divisions) against the ~7 us round trip it replaces -- align 0.257 ms against 0.216.
 *   "device_newton_fault_at" (default 0)  test hook: n > 0 makes the n-th comparison of every align fail. */
int ndtb_set_option(ndtb_handle h, const char* name, double value);
/* Life-time counters of the device Newton step: stats[0] steps taken on the device and confirmed by the host, stats[1] steps
 * the device handed to the host, stats[2] comparisons that failed (each closes the resident kernel of that align), stats[3] 1 when
 * the library's restatement of the host libm's sincosf (glibc >= 2.28, x86-64 FMA build) reproduces the running libm bit for bit
 * on a sample of 24 000 arguments -- with 0 the device step is never used, whatever "device_newton" says.  h may be NULL (the
 * counters then read 0; stats[3] needs no GPU). */
int ndtb_get_session_stats(ndtb_handle h, unsigned long long stats[4]);
/* The Newton-direction solve of ndtb_align on its own (host code, no GPU): x = H^-1 b by the path "direct_solver"
 * selects (method 1; falls back to the SVD when H is not certified well conditioned) or by the Jacobi SVD with
 * Eigen's rank threshold (method 0).  *used_direct reports which one produced x.  Parity hook for the tests. */
/* The step after align in the reference's pipelines (run/pipeline_ligo_tc.cpp:536-542, src/registercallback.cpp:170-186):
 * lidar factor covariance = -(hessian + 1e-6 I)^-1, reordered from NDT's (translation, rotation) to GTSAM's
 * (rotation, translation) block order.  Host code, row-major 6x6 in and out; a singular matrix yields NaNs (as
 * Eigen's inverse() would yield non-finite values). */
int ndtb_hessian_to_covariance(const double hessian[36], double covariance_gtsam[36]);
int ndtb_newton_direction(const double H[36], const double b[6], int method, double x[6], int* used_direct);
/* Duration of ONE derivative evaluation kernel (DIRECT1/7 as configured, with Hessian) at pose p / transform T16,
 * measured as `reps` back-to-back launches between a single pair of CUDA events on the handle's stream: the event
 * overhead (~7 us per pair on a B200 box) is amortised over the launches instead of being added to each. */
int ndtb_time_eval_kernel(ndtb_handle h, const double p[6], const float* T16, int reps, float* mean_us);
/* Test hook: ONE Newton step between two evaluations of ndtb_align (direction by pivoted elimination, normalisation, descent flip,
 * step clamp, pose -> float 4x4, angle tables; ndt_omp_impl.hpp:128-190, 396-538, 915-990), computed by the warp-cooperative device
 * code (which = 0, device_newton.cuh) or by the host code the align loop runs (which = 1), for bit-for-bit comparison.
 * params = {step_max, step_min, trans_epsilon, -}; status: 0 continue (out_state81 = T[12], j_ang[24], h_ang[45]; out_p6 = the next
 * pose), 1 converged, 2 the host decides (ill-conditioned H -> SVD, zero / NaN step, zero slope). */
int ndtb_test_newton_step(int which, int cuda_device, const double g[6], const double H[36], const double p[6], const double params[4],
                          int max_iterations, int first, int nr_iterations, double a_t_prev, int* status, int* nr_out, double* a_t_out,
                          float* out_state81, double* out_p6);
/* Number of CUDA kernels this handle has launched since creation (bench.py's gpu_launches). */
int ndtb_get_launch_count(ndtb_handle h, uint64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* NDTB_H_ */
