"""Host-side mirror of the reference's registration classes over the C ABI (same method names,
argument meaning and error behaviour), so parity tests read like the reference's own tests
(extern/svn_ndt/test/test_svn_ndt.cpp:138-317).

  NormalDistributionsTransform      <- pclomp::NormalDistributionsTransform (ndt_omp.h:78-553)
  SvnNormalDistributionsTransform   <- svn_ndt::SvnNormalDistributionsTransform (svn_ndt.h:65-316)

The C++ callers (src/registercallback.cpp, run/pipeline_*.cpp) use include/ndtb_pclomp.hpp instead;
both sit on the same libndtb.so entry points.
"""
import numpy as np

from . import ndtb

KDTREE, DIRECT26, DIRECT7, DIRECT1 = ndtb.KDTREE, ndtb.DIRECT26, ndtb.DIRECT7, ndtb.DIRECT1


class NdtResult:
    """pclomp::NdtResult (ndt_struct.hpp:67-88)."""

    def __init__(self, d):
        self.pose = d["T"]
        self.transform_probability = d["trans_probability"]
        self.nearest_voxel_transformation_likelihood = d["nvtl"]
        self.iteration_num = d["iterations"]
        self.transformation_array = list(d["transformation_array"])
        self.hessian = d["hessian"]


class NormalDistributionsTransform:
    def __init__(self, device=0):
        self._h = ndtb.Handle(device)
        self._h.set_params(grid_flavour=ndtb.FLAVOUR_PCLOMP)
        self._target = None
        self._source = None
        self._last = None

    # ---- configuration (ndt_omp.h:119-291) ----
    def setNumThreads(self, n):  # noqa: N802  (kept for call-site compatibility; the GPU has no thread knob)
        self._num_threads = n

    def getNumThreads(self):  # noqa: N802
        return getattr(self, "_num_threads", 0)

    def setResolution(self, r):  # noqa: N802
        self._check(self._h.set_params(resolution=float(r)), allow_empty=True)

    def getResolution(self):  # noqa: N802
        return self._h.get_params().resolution

    def setStepSize(self, s):  # noqa: N802
        self._h.set_params(step_size=float(s))

    def getStepSize(self):  # noqa: N802
        return self._h.get_params().step_size

    def setTransformationEpsilon(self, e):  # noqa: N802
        self._h.set_params(trans_epsilon=float(e))

    def setMaximumIterations(self, n):  # noqa: N802
        self._h.set_params(max_iterations=int(n))

    def setNeighborhoodSearchMethod(self, m):  # noqa: N802
        rc = self._h.set_params(search_method=int(m))
        if rc != ndtb.NDTB_OK:
            raise ndtb.NdtbError(rc, "KDTREE / DIRECT26 / DIRECT7 / DIRECT1 are provided (the first two without line search)")

    def setOulierRatio(self, r):  # noqa: N802  (sic: the reference spells it this way, ndt_omp.h:173)
        self._h.set_params(outlier_ratio=float(r))

    def setRegularizationScaleFactor(self, f):  # noqa: N802
        self._h.set_params(regularization_scale_factor=float(f))

    def setRegularizationPose(self, T):  # noqa: N802
        self._h.set_regularization_pose(T)

    def unsetRegularizationPose(self):  # noqa: N802
        self._h.set_regularization_pose(None)

    def setMinPointPerVoxel(self, n):  # noqa: N802
        self._check(self._h.set_params(min_points_per_voxel=int(n)), allow_empty=True)

    # ---- clouds ----
    def setInputTarget(self, cloud):  # noqa: N802
        self._target = cloud
        rc = self._h.set_target(cloud)
        # the reference only warns on an empty / overflowing target and leaves the grid empty
        if rc not in (ndtb.NDTB_OK, ndtb.NDTB_ERR_EMPTY, ndtb.NDTB_ERR_OVERFLOW):
            self._h.check(rc)
        self.last_status = rc

    def setInputSource(self, cloud):  # noqa: N802
        self._source = cloud
        rc = self._h.set_source(cloud)
        if rc not in (ndtb.NDTB_OK, ndtb.NDTB_ERR_EMPTY):
            self._h.check(rc)

    # ---- registration ----
    def align(self, guess=None):
        self._last = self._h.align(guess)
        return self._last

    def computeTransformation(self, guess=None):  # noqa: N802  (test_svn_ndt.cpp:171 calls it directly)
        return self.align(guess)

    def hasConverged(self):  # noqa: N802
        return bool(self._last and self._last["converged"])

    def getFinalNumIteration(self):  # noqa: N802
        return self._last["iterations"]

    def getFinalTransformation(self):  # noqa: N802
        return self._last["T"]

    def getTransformationProbability(self):  # noqa: N802
        return self._last["trans_probability"]

    def getNearestVoxelTransformationLikelihood(self):  # noqa: N802
        return self._last["nvtl"]

    def getHessian(self):  # noqa: N802
        return self._last["hessian"]

    def getFinalTransformationArray(self):  # noqa: N802
        return list(self._last["transformation_array"])

    def getResult(self):  # noqa: N802
        return NdtResult(self._last)

    def calculateTransformationProbability(self, T):  # noqa: N802
        return self._h.calculate_scores(T)[0]

    def calculateNearestVoxelTransformationLikelihood(self, T):  # noqa: N802
        return self._h.calculate_scores(T)[1]

    def getTargetCells(self):  # noqa: N802
        return self._h.voxels()

    # parity hook (protected computeDerivatives in the reference)
    def computeDerivatives(self, p, T=None, compute_hessian=True):  # noqa: N802
        return self._h.compute_derivatives(p, T, compute_hessian)

    def _check(self, rc, allow_empty=False):
        if rc == ndtb.NDTB_OK or (allow_empty and rc in (ndtb.NDTB_ERR_EMPTY, ndtb.NDTB_ERR_OVERFLOW)):
            return
        self._h.check(rc)

    @property
    def handle(self):
        return self._h


class MultiGridNormalDistributionsTransform(NormalDistributionsTransform):
    """pclomp::MultiGridNormalDistributionsTransform (extern/ndt_omp/include/multigrid_pclomp/multigrid_ndt_omp.h):
    the target is a set of voxel grids keyed by an id (addTarget / removeTarget / createVoxelKdtree), the only
    neighbour search is the radius search over leaf centroids, derivatives are FP64.  Setters are the base class's."""

    def setNeighborhoodSearchMethod(self, m):  # noqa: N802  multigrid_ndt_omp_impl.hpp:197: KDTREE only
        if int(m) != KDTREE:
            raise ndtb.NdtbError(ndtb.NDTB_ERR_INVALID, "only KDTREE is supported in multigrid_ndt_omp")

    def setInputTarget(self, cloud):  # noqa: N802  multigrid_ndt_omp.h:143-148
        self.addTarget(cloud, "default")
        self.createVoxelKdtree()

    def addTarget(self, cloud, target_id):  # noqa: N802  multigrid_ndt_omp.h:150-155
        rc = self._h.mg_add_target(cloud, target_id)
        if rc not in (ndtb.NDTB_OK, ndtb.NDTB_ERR_EMPTY, ndtb.NDTB_ERR_OVERFLOW):   # the reference only warns on those
            self._h.check(rc)
        self.last_status = rc

    def removeTarget(self, target_id):  # noqa: N802  multigrid_ndt_omp.h:157
        self._h.check(self._h.mg_remove_target(target_id))

    def createVoxelKdtree(self):  # noqa: N802  multigrid_ndt_omp.h:159-163
        self._h.check(self._h.mg_create_kdtree())

    def getCurrentMapIDs(self):  # noqa: N802  multigrid_ndt_omp.h:325
        return self._h.mg_grid_ids()

    def getVoxelPCD(self):  # noqa: N802  multigrid_ndt_omp.h:323: the leaf centroids the kd-tree is built over
        return self._h.mg_leaves()["centroid"]

    def align(self, guess=None):
        self._last = self._h.mg_align(guess)
        return self._last

    def calculateTransformationProbability(self, T):  # noqa: N802
        return self._h.mg_calculate_scores(T)[0]

    def calculateNearestVoxelTransformationLikelihood(self, T):  # noqa: N802
        return self._h.mg_calculate_scores(T)[1]

    def getTargetCells(self):  # noqa: N802
        return self._h.mg_leaves()

    def computeDerivatives(self, p, T=None, compute_hessian=True):  # noqa: N802
        return self._h.mg_compute_derivatives(p, T, compute_hessian)


class SvnNdtResult:
    """svn_ndt::SvnNdtResult (svn_ndt.h:40-51)."""

    def __init__(self, final_pose, final_covariance, converged, iterations, particles=None):
        self.final_pose = final_pose
        self.final_covariance = final_covariance
        self.converged = converged
        self.iterations = iterations
        self.particles = particles


class SvnNormalDistributionsTransform:
    def __init__(self, device=0):
        self._h = ndtb.Handle(device)
        self._h.set_params(grid_flavour=ndtb.FLAVOUR_SVN)
        self._h.set_svn_params()
        self._target = None

    def setResolution(self, r):  # noqa: N802  svn_ndt_impl.hpp:162-176
        if r <= 1e-6:
            return
        self._h.set_params(resolution=float(r))

    def setMinPointPerVoxel(self, n):  # noqa: N802  svn_ndt_impl.hpp:179-187 (clamped to >= 3)
        self._h.set_params(min_points_per_voxel=max(3, int(n)))

    def setNeighborhoodSearchMethod(self, m):  # noqa: N802
        rc = self._h.set_params(search_method=int(m))
        if rc != ndtb.NDTB_OK:
            raise ndtb.NdtbError(rc, "KDTREE / DIRECT7 / DIRECT1 are provided (svn_ndt.h:31-36)")

    def setNumThreads(self, n):  # noqa: N802
        self._num_threads = n

    def getNumThreads(self):  # noqa: N802
        return getattr(self, "_num_threads", 0)

    def setParticleCount(self, k):  # noqa: N802
        self._h.set_svn_params(particle_count=int(k))

    def setMaxIterations(self, n):  # noqa: N802
        self._h.set_svn_params(max_iterations=int(n))

    def setKernelBandwidth(self, h):  # noqa: N802
        self._h.set_svn_params(kernel_bandwidth=float(h))

    def setStepSize(self, s):  # noqa: N802
        self._h.set_svn_params(step_size=float(s))

    def setEarlyStopThreshold(self, t):  # noqa: N802
        self._h.set_svn_params(stop_threshold=float(t))

    def setUseGaussNewtonHessian(self, b):  # noqa: N802
        self._h.set_svn_params(use_gauss_newton=int(bool(b)))

    def setOutlierRatio(self, r):  # noqa: N802  svn_ndt_impl.hpp:190-207 (clamped to [0, 1))
        r = 0.0 if r < 0.0 else (1.0 - 1e-9 if r >= 1.0 else r)
        self._h.set_params(outlier_ratio=float(r))

    def setInputTarget(self, cloud):  # noqa: N802
        rc = self._h.set_target(cloud)
        if rc not in (ndtb.NDTB_OK, ndtb.NDTB_ERR_EMPTY, ndtb.NDTB_ERR_OVERFLOW):
            self._h.check(rc)

    def align(self, source_cloud, prior_mean, init_particles=None, seed=0):
        """align(source, prior) -> SvnNdtResult.  On invalid inputs the reference logs an error and
        returns converged=False, final_pose=prior, final_covariance=I (svn_ndt_impl.hpp:682-702)."""
        prior = np.ascontiguousarray(prior_mean, np.float64).reshape(4, 4)
        rc = self._h.set_source(source_cloud) if source_cloud is not None and len(source_cloud) else ndtb.NDTB_ERR_EMPTY
        if rc != ndtb.NDTB_OK:
            return SvnNdtResult(prior.copy(), np.eye(6), False, 0)
        try:
            r = self._h.svn_align(prior, init_particles, seed)
        except ndtb.NdtbError as e:
            if e.code in (ndtb.NDTB_ERR_NOT_READY, ndtb.NDTB_ERR_EMPTY, ndtb.NDTB_ERR_INVALID):
                return SvnNdtResult(prior.copy(), np.eye(6), False, 0)
            raise
        return SvnNdtResult(r["final_pose"], r["final_covariance"], r["converged"], r["iterations"], r["particles"])

    @property
    def handle(self):
        return self._h
