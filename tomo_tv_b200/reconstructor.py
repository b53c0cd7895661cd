"""`TomoGPU` -- the user-facing reconstruction object, API-compatible with the reference's
tomofusion.gpu.reconstructor.TomoGPU (/root/reference/tomofusion/gpu/reconstructor.py:12-215):

    TomoGPU(tiltAngles[deg], tiltSeries[Nslice, Nray, Nangles], gpu_id=-1)
    .wbp(filter) .sart(Niter, init) .sirt(Niter) .cgls(Niter) .kl_divergence(Niter, lambda_param)
    .fista(Niter, momentum, lambda_param, nTViter, show_convergence)
    .asd_pocs(Niter, eps, beta0, beta_reduce, r_max, nTViter, alpha, alpha_reduce)
    .get_recon() -> float64 [Nslice, Nray, Nray]

Behavioural notes, all kept on purpose:
  * `fista` with momentum runs tv_fgp on `recon` and then `fista_momentum` overwrites `recon`
    with `yk` (reference loop, reconstructor.py:137-148) -- bug-compatible.
  * `_run_iterative` always evaluates the data distance (the caller's show_convergence flag is
    ignored in the reference, reconstructor.py:100); here the flag is honoured for the *plot*
    only, the cost vector is still filled.
  * the reference's `asd_pocs` cannot run (undefined names, no SART initialisation,
    reconstructor.py:157-192); this one implements the canonical loop of
    tomofusion/examples/sim_ASD.py:66-94 behind the same signature.

The tilt series is packed on the host with one vectorised transpose (the reference loops over
slices in Python, reconstructor.py:54-56) and the volume is fetched with one bulk copy.
Plotting (matplotlib) is imported lazily and skipped when unavailable; the Tk slice viewer of the
reference is out of scope.
"""
import numpy as np

from . import pytvlib
from .engine import tomoengine, device_count


def _progress(n):
    try:
        from tqdm import tqdm
        return tqdm(range(n))
    except Exception:  # pragma: no cover
        return range(n)


def pack_tilt_series(tiltSeries):
    """[Nslice, Nray, Nangles] -> b[s, a*Nray + d] = tiltSeries[s, d, a]  (reconstructor.py:54-56)."""
    ts = np.asarray(tiltSeries)
    if ts.ndim != 3:
        raise ValueError("tiltSeries must be a 3-D array [Nslice, Nray, Nangles]")
    ns, nr, na = ts.shape
    return np.ascontiguousarray(np.transpose(ts, (0, 2, 1)).reshape(ns, na * nr), dtype=np.float32)


class TomoGPU:

    def __init__(self, tiltAngles, tiltSeries=None, gpu_id: int = -1, verbose: bool = True):
        pytvlib.check_cuda()
        if tiltSeries is None:
            raise ValueError("TomoGPU needs a tilt series [Nslice, Nray, Nangles]")
        self.verbose = verbose
        self.tiltAngles = np.asarray(tiltAngles, dtype=np.float64)
        self.Nslice, self.Nray, self.Nangles = np.shape(tiltSeries)
        if self.tiltAngles.size != self.Nangles:
            raise ValueError("number of tilt angles does not match tiltSeries.shape[2]")
        # One engine drives one GPU.  The reference's 'multigpu' configuration (gpu_id < 0 and several
        # GPUs, tomofusion/__init__.py:21-34) is one process per GPU here: under torchrun every rank
        # builds the same TomoGPU on the full tilt series and owns a contiguous x-slab of it
        # (tomo_tv_b200.sharded.multigpuengine); get_recon() returns the whole volume on every rank.
        ngpu = device_count()
        if ngpu == 0:
            raise ValueError('An Nvidia GPU is Needed for this package!')
        self._dist = None
        if gpu_id < 0:
            try:
                import torch.distributed as td
                if td.is_available() and td.is_initialized() and td.get_world_size() > 1:
                    self._dist = td
            except ImportError:
                pass
        if self._dist is not None:
            from .sharded import multigpuengine
            self._say("Initializing multi-GPU configuration")
            self.tomo = multigpuengine(self.Nslice, self.Nray, np.deg2rad(self.tiltAngles), dist=self._dist)
        else:
            self._say("Initializing single-GPU configuration")
            self.tomo = tomoengine(self.Nslice, self.Nray, np.deg2rad(self.tiltAngles),
                                   device=gpu_id if gpu_id >= 0 else -1)
            if gpu_id >= 0:
                self.tomo.set_gpu(gpu_id)
        self.set_tilt_series(tiltSeries)
        self.recon = None
        self.cost = None

    # ------------------------------------------------------------------------------------------
    def set_tilt_series(self, tiltSeries):
        shp = np.shape(tiltSeries)
        if tuple(shp) != (self.Nslice, self.Nray, self.Nangles):
            raise ValueError("tilt series shape changed; build a new TomoGPU")
        self.recon = None
        b = pack_tilt_series(tiltSeries)
        if self._dist is not None:
            from .sharded import slab_of
            b = slab_of(b, self._dist)
        self.tomo.set_tilt_series(b)

    def _say(self, msg):
        if self.verbose:
            print(msg)

    # ------------------------------------------------------------------------------------------
    def wbp(self, filter: str = 'ram-lak'):
        if filter not in pytvlib.wbp_filters():
            print(f'{filter} Filter not supported. Defaulting to ram-lak.')
            filter = 'ram-lak'
        pytvlib.initialize_algorithm(self.tomo, 'FBP', filter)
        pytvlib.run(self.tomo, 'FBP')

    def sart(self, Niter: int = 150, init: str = 'sequential', show_convergence: bool = True):
        if init not in pytvlib.sart_orders():
            print(f'{init} initialization method not supported. Defaulting to sequential.')
            init = 'sequential'
        pytvlib.initialize_algorithm(self.tomo, 'SART', init)
        self._run_iterative('SART', Niter, show_convergence)

    def sirt(self, Niter: int = 150, show_convergence: bool = True):
        pytvlib.initialize_algorithm(self.tomo, 'SIRT')
        self._run_iterative('SIRT', Niter, show_convergence)

    def cgls(self, Niter: int = 100, show_convergence: bool = True):
        pytvlib.initialize_algorithm(self.tomo, 'CGLS')
        self._run_iterative('CGLS', Niter, show_convergence)

    def _run_iterative(self, alg, Niter, show_convergence=True):
        self.cost = np.zeros(Niter)
        self.tomo.restart_recon()
        for i in _progress(Niter):
            pytvlib.run(self.tomo, alg)
            self.cost[i] = self.tomo.data_distance()
        if show_convergence:
            self.plot_convergence(self.cost, alg)

    def kl_divergence(self, Niter: int = 100, lambda_param: float = 0.1):
        self.tomo.restart_recon()
        self._say('Running Reconstruction...')
        pytvlib.initialize_algorithm(self.tomo, 'kl-divergence')
        self.cost = np.zeros(Niter)
        for i in _progress(Niter):
            self.cost[i] = pytvlib.run(self.tomo, 'kl-divergence', lambda_param)

    def fista(self, Niter: int = 100, momentum: bool = True, lambda_param: float = 0.1,
              nTViter: int = 10, show_convergence: bool = True):
        pytvlib.initialize_algorithm(self.tomo, 'fista')
        if not momentum:
            self.tomo.remove_momentum()
        self.cost = np.zeros(Niter)
        t_prev = 1.0
        self._say('Running Reconstruction...')
        for k in _progress(Niter):
            pytvlib.run(self.tomo, 'fista')                 # one SIRT step (pytvlib.py:27)
            self.tomo.tv_fgp(nTViter, lambda_param)
            if momentum:
                t_next = 0.5 * (1.0 + np.sqrt(1.0 + 4.0 * t_prev ** 2))
                self.tomo.fista_momentum((t_prev - 1.0) / t_next)
                t_prev = t_next
            if show_convergence:
                self.cost[k] = 0.5 * self.tomo.data_distance() ** 2 + lambda_param * self.tomo.tv()
        if show_convergence:
            self.plot_convergence(self.cost, 'FISTA')

    def asd_pocs(self, Niter: int = 100, eps: float = 0.025, beta0: float = 0.25,
                 beta_reduce: float = 0.9985, r_max: float = 0.95, nTViter: int = 10,
                 alpha: float = 0.2, alpha_reduce: float = 0.95, show_convergence: bool = True):
        # canonical loop: tomofusion/examples/sim_ASD.py:66-94
        pytvlib.initialize_algorithm(self.tomo, 'asd-pocs', 'sequential')
        self.tomo.initialize_recon_copy()
        self.tomo.restart_recon()
        beta = float(beta0)
        self.dd_vec, self.tv_vec = np.zeros(Niter), np.zeros(Niter)
        dPOCS = 0.0
        self._say('Running Reconstruction...')
        for i in _progress(Niter):
            self.tomo.copy_recon()
            pytvlib.run(self.tomo, 'sart', beta)
            beta *= beta_reduce
            dp = self.tomo.matrix_2norm()
            if i == 0:
                dPOCS = dp * alpha
            self.dd_vec[i] = self.tomo.data_distance()
            self.tomo.copy_recon()
            self.tv_vec[i] = self.tomo.tv_gd(nTViter, dPOCS)
            dg = self.tomo.matrix_2norm()
            if dg > dp * r_max and self.dd_vec[i] > eps:
                dPOCS *= alpha_reduce
        self.cost = self.dd_vec

    # ------------------------------------------------------------------------------------------
    def plot_convergence(self, cost, algorithm):
        try:
            import matplotlib.pyplot as plt
        except Exception:
            return                                          # headless box: nothing to draw on
        n = cost.shape[0]
        plt.figure(figsize=(8, 5))
        plt.scatter(np.arange(n), cost)
        plt.xlabel('Iteration'); plt.ylabel('Cost')
        plt.title(f'{algorithm} Convergence')
        plt.xlim([0, n - 1])
        plt.show()

    def get_recon(self):
        """float64 [Nslice, Nray, Nray] like the reference (reconstructor.py:207-215), one bulk copy."""
        rec = self.tomo.get_recon_all()
        if self._dist is not None:
            from .sharded import gather_slabs
            rec = gather_slabs(rec, self._dist)
        self.recon = rec.astype(np.float64)
        return self.recon

    def get_projections(self):
        """The measured tilt series as the constructor takes it, float32 [Nslice, Nray, Nangles]
        (a `pass` stub in the reference, reconstructor.py:217-219; with several ranks: this rank's slab)."""
        b = self.tomo.get_projections()
        return np.ascontiguousarray(b.reshape(b.shape[0], self.Nangles, self.Nray).transpose(0, 2, 1))

    def show_recon(self):
        """The reference's Tkinter slice viewer (reconstructor.py:221-383) is UI, outside the scope of this
        package: fetch the volume with get_recon() and view it with the tool of your choice."""
        raise RuntimeError("show_recon: the Tk viewer is not part of tomo_tv_b200; use get_recon()")

