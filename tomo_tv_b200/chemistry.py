"""`ChemicalTomo` -- API-compatible with the reference's tomofusion.chemistry.reconstructor.ChemicalTomo
(/root/reference/tomofusion/chemistry/reconstructor.py:13-243) on top of the CUDA fusion engine
(`tomo_tv_b200.fusion.multimodal`).

    ChemicalTomo(haadf[Nx,Ny,NaH], haadfTiltAngles[deg], chem{el: [Nx,Ny,NaC]}, chemTiltAngles[deg],
                 gamma=1.6, sigmaMethod=3, gpu_id=-1)
    .chemical_tomography(Niter, lambdaCHEM, alg, show_convergence)
    .data_fusion(Niter, lambdaCHEM, lambdaHAADF, lambdaTV, iterSIRT, tvIter, show_convergence)
    .get_recon() -> float32 [Nel, Nx, Ny, Ny]

Kept from the reference on purpose: the inputs are normalised IN PLACE (reconstructor.py:102-103,
:119-120); the summation matrix is always built with gamma 1.6 and float16-rounded weights
(reconstructor.py:147, fusion_helper.py:6); lambdaCHEM shrinks by 0.95 whenever the HAADF cost rises
(:203).  The Tk viewer and matplotlib plots are UI and out of scope (plots are drawn only when
matplotlib is importable).
"""
import numpy as np

from .fusion import multimodal

_PT = ("h he li be b c n o f ne na mg al si p s cl ar k ca sc ti v cr mn fe co ni cu zn ga ge as se br kr rb sr y zr nb mo "
       "tc ru rh pd ag cd in sn sb te i xe cs ba la ce pr nd pm sm eu gd tb dy ho er tm yb lu hf ta w re os ir pt au hg tl "
       "pb bi po at rn fr ra ac th pa u np pu am cm bk cf es fm md no lr rf").split()


def get_periodic_table():
    """element symbol (lower case) -> atomic number  (utils/fusion_helper.py:36-52)."""
    return {s: i + 1 for i, s in enumerate(_PT)}


def summation_weights(zNums, gamma, method=0):
    """Per-element weights of the summation matrix, float16-rounded like the reference's `vals`
    array (utils/fusion_helper.py:6,18-27)."""
    z = np.asarray(zNums, dtype=np.float64)
    if method == 0:
        w = np.ones_like(z)
    elif method == 1:
        w = z / np.mean(z)
    elif method == 2:
        w = z ** gamma / np.mean(z ** gamma)
    elif method == 3:
        w = z / np.sum(z)
    elif method == 4:
        w = z ** gamma / np.sum(z ** gamma)
    else:
        w = np.ones_like(z)
    return w.astype(np.float16).astype(np.float32)


def create_weighted_summation_matrix(nx, ny, nz, zNums, gamma, method=0):
    """Same matrix as utils/fusion_helper.py:5-32, returned directly in the [3, nnz] float32
    rows / cols / vals form `load_sigma` takes (entry (p, p + nx*ny*e) = w_e), built without the
    reference's Python loop over pixels."""
    w = summation_weights(zNums, gamma, method)
    npix = nx * ny
    p = np.arange(npix, dtype=np.int64)
    rows = np.repeat(p, nz)
    cols = (p[:, None] + npix * np.arange(nz, dtype=np.int64)[None, :]).ravel()
    vals = np.tile(w, npix)
    return np.array([rows, cols, vals], dtype=np.float32, order='C')


def _progress(n):
    try:
        from tqdm import tqdm
        return tqdm(range(n))
    except Exception:  # pragma: no cover
        return range(n)


class ChemicalTomo:

    def __init__(self, haadf, haadfTiltAngles, chem, chemTiltAngles, gamma: float = 1.6, sigmaMethod: int = 3,
                 gpu_id: int = -1, verbose: bool = True):
        (self.nx, self.ny, _) = haadf.shape
        self.elements = list(chem)
        self.nz = len(chem)
        self.verbose = verbose
        haadfTiltAngles = np.asarray(haadfTiltAngles, dtype=np.float64)
        chemTiltAngles = np.asarray(chemTiltAngles, dtype=np.float64)
        self.NprojHAADF = haadfTiltAngles.shape[0]
        self.NprojCHEM = chemTiltAngles.shape[0]
        # One engine drives one GPU.  The reference's 'multigpu' configuration (reconstructor.py:53-63: multigpufusion fans
        # slices out over the GPUs with OpenMP) is one process per GPU here: under torchrun every rank builds the same
        # ChemicalTomo on the full tilt series and owns a contiguous x-slab (fusion.multigpufusion); get_recon() returns the
        # whole 4-D volume on every rank.
        self._dist = None
        if gpu_id < 0:
            try:
                import torch.distributed as td
                if td.is_available() and td.is_initialized() and td.get_world_size() > 1:
                    self._dist = td
            except ImportError:
                pass
        if self._dist is not None:
            from .fusion import multigpufusion
            self.tomo = multigpufusion(self.nx, self.ny, self.nz, np.deg2rad(haadfTiltAngles), np.deg2rad(chemTiltAngles), dist=self._dist)
        else:
            self.tomo = multimodal(self.nx, self.ny, self.nz, np.deg2rad(haadfTiltAngles), np.deg2rad(chemTiltAngles),
                                   device=gpu_id if gpu_id >= 0 else -1)
            if gpu_id >= 0:
                self.tomo.set_gpu(gpu_id)

        self.set_haadf_projections(haadf)
        self.set_chemical_projections(chem)
        self.set_summation_matrix(gamma, sigmaMethod)

        self.gamma = gamma
        self.sigmaMethod = sigmaMethod
        self.reduceLambda = True

        self.tomo.initialize_FP(); self.tomo.initialize_BP(); self.tomo.initialize_SIRT()
        self.tomo.estimate_lipschitz()
        self.tomo.set_measureChem(True); self.tomo.set_measureHaadf(True)

        self.reconTotal = None
        self.chemistry_reconstructed = False
        self.costs = {}

    def set_haadf_projections(self, haadf):
        """reconstructor.py:96-109 (normalises `haadf` in place, like the reference)."""
        haadf[haadf < 0] = 0
        haadf /= np.max(haadf)
        bh = np.ascontiguousarray(np.transpose(haadf, (0, 2, 1)).reshape(self.nx, self.ny * self.NprojHAADF), dtype=np.float32)
        self.tomo.set_haadf_tilt_series(self._slab(bh))

    def set_chemical_projections(self, chem):
        """reconstructor.py:111-137: bChem[s] = concat_e chem[e][s].T.flatten()."""
        for element in self.elements:
            chem[element][chem[element] < 0] = 0
            chem[element] /= np.max(chem[element])
        parts = [np.transpose(chem[el], (0, 2, 1)).reshape(self.nx, self.ny * self.NprojCHEM) for el in self.elements]
        bChem = np.ascontiguousarray(np.concatenate(parts, axis=1), dtype=np.float32)
        self.tomo.set_chem_tilt_series(self._slab(bChem))

    def _slab(self, rows):
        """This rank's rows of a [Nslice, ...] host array (the whole array on one GPU)."""
        if self._dist is None:
            return rows
        from .sharded import slab_of
        return slab_of(rows, self._dist)

    def set_summation_matrix(self, gamma: float = 1.6, sigmaMethod: int = 3):
        """reconstructor.py:139-153."""
        self.tomo.set_gamma(gamma)
        pt_table = get_periodic_table()
        zNums = np.array([pt_table[x.lower()] for x in self.elements])
        sigma = create_weighted_summation_matrix(self.ny, self.ny, self.nz, zNums, 1.6, sigmaMethod)
        self.tomo.load_sigma(sigma)

    def chemical_tomography(self, Niter: int = 100, lambdaCHEM: float = 0.05, alg='kldivergnce', show_convergence: bool = True):
        """reconstructor.py:155-181."""
        self.tomo.restart_recon()
        costCHEM = np.zeros(Niter)
        for ii in (_progress(Niter) if self.verbose else range(Niter)):
            costCHEM[ii] = self.tomo.poisson_ml(lambdaCHEM)
        self.costs['chemical_tomography'] = costCHEM
        if show_convergence:
            self._plot([costCHEM], ['Cost'])
        self.chemistry_reconstructed = True

    def data_fusion(self, Niter: int = 50, lambdaCHEM: float = 5e-2, lambdaHAADF: float = 10, lambdaTV: float = 1e-4,
                    iterSIRT: int = 5, tvIter: int = 5, show_convergence: bool = True):
        """reconstructor.py:182-225."""
        if not self.chemistry_reconstructed:
            self.chemical_tomography(lambdaCHEM=lambdaCHEM, show_convergence=show_convergence)
        self._rescale_data()
        costCHEM = np.zeros(Niter, dtype=np.float32)
        costHAADF = costCHEM.copy(); costTV = costCHEM.copy()
        for i in (_progress(Niter) if self.verbose else range(Niter)):
            (costHAADF[i], costCHEM[i]) = self.tomo.sirt_data_fusion(lambdaHAADF, lambdaCHEM, iterSIRT)
            costTV[i] = self.tomo.tv_fgp_4D(tvIter, lambdaTV)
            if i > 0 and costHAADF[i] > costHAADF[i - 1]:
                lambdaCHEM *= 0.95
        self.costs['data_fusion'] = (costHAADF, costCHEM, costTV)
        if show_convergence:
            self._plot([costHAADF, costCHEM, costTV], [r'$||A (\Sigma x) - b||^2$', r'$\sum (Ax - b \cdot \log(Ax))$', r'$\sum \|x\|_{TV}$'])

    def _rescale_data(self, scale: float = 10):
        """reconstructor.py:227-236."""
        self.tomo.rescale_tomograms(scale)
        self.tomo.rescale_projections()

    def get_recon(self):
        """reconstructor.py:238-249 with one bulk copy per element instead of Nel*Nx slice calls."""
        if self.reconTotal is None:
            self.reconTotal = np.zeros([self.nz, self.nx, self.ny, self.ny], dtype=np.float32)
        for e in range(self.nz):
            rec = self.tomo.get_recon_all(e)
            if self._dist is not None:
                from .sharded import gather_slabs
                rec = gather_slabs(rec, self._dist)
            self.reconTotal[e] = rec
        return self.reconTotal

    @staticmethod
    def _plot(series, labels):
        try:
            import matplotlib.pyplot as plt
        except Exception:
            return
        plt.figure(figsize=(9, 2 * len(series) + 2))
        for i, (s, l) in enumerate(zip(series, labels)):
            ax = plt.subplot(len(series), 1, i + 1)
            ax.plot(s); ax.set_ylabel(l); ax.set_xlim([0, max(len(s) - 1, 1)])
        plt.xlabel('# Iterations')
        plt.show()

    def display_recon(self):
        """The reference's Tkinter element viewer (chemistry/reconstructor.py:251-...) is UI, outside the scope
        of this package: fetch the 4-D array with get_recon()."""
        raise RuntimeError("display_recon: the Tk viewer is not part of tomo_tv_b200; use get_recon()")
