"""Algorithm-name dispatch, same contract as the reference's tomofusion/pytvlib.py:5-39.

The dispatch helpers, plus the reference's result files (pytvlib.py:120-162: `/parameters` attributes,
`/results/<name>` float32 datasets, `/Reconstruction/recon` with Nslice / Nray / Nproj attributes).  With h5py the
files are the reference's HDF5 layout; without it (this image has no h5py) the same tree is written as an `.npz`
whose keys are the HDF5 paths (`results/cost`, `parameters@lambda`, `Reconstruction/recon`, `Reconstruction@Nslice`).
TIFF / skimage loading (pytvlib.py:57-118) is not reproduced.
"""
import os

import numpy as np

_SIRT = ('SIRT', 'sirt')
_CGLS = ('CGLS', 'cgls')
_FISTA = ('FISTA', 'fista')
_KL = ('poisson_ML', 'kl-divergence')
_SART = ('SART', 'sart')
_ASD = ('asd-pocs', 'ASD-POCS')
_FBP = ('FBP', 'WBP', 'fbp', 'wbp')


def initialize_algorithm(tomo, alg, initAlg=''):
    """Prepare `tomo` for `alg`; always finishes with initialize_FP (pytvlib.py:5-19)."""
    if alg in _SIRT:
        tomo.initialize_SIRT()
    elif alg in _CGLS:
        tomo.initialize_CGLS()
    elif alg in _FISTA:
        tomo.initialize_fista()
    elif alg in _KL:
        tomo.initialize_poisson_ML()
    elif alg in _SART or alg in _ASD:
        tomo.initialize_SART(initAlg)          # initAlg in {'sequential', 'random'}
    elif alg in _FBP:
        tomo.initialize_FBP(initAlg)           # initAlg = filter name
    tomo.initialize_FP()


def run(tomo, alg, beta=1, niter=1):
    """One outer step of `alg` (pytvlib.py:21-31).  Note 'fista' is one SIRT iteration."""
    if alg in _SIRT or alg in _FISTA:
        tomo.SIRT(niter)
    elif alg in _CGLS:
        tomo.CGLS(niter)
    elif alg in _SART:
        tomo.SART(beta, niter)
    elif alg in _FBP:
        tomo.FBP(True)
    elif alg in _KL:
        return tomo.poisson_ML(beta)
    return None


def wbp_filters():
    return ['ram-lak', 'shepp-logan', 'hamming', 'cosine', 'parzen', 'lanczos',
            'triangular', 'gaussian', 'blackman', 'nuttall', 'blackman-harris', 'kaiser']


def sart_orders():
    return ['sequential', 'random']


def check_cuda():
    """The reference probes ASTRA (pytvlib.py:42-51); here: is a CUDA device usable at all."""
    from .engine import device_count
    if device_count() == 0:
        print('No GPU support available')
        print('Please have a GPU to run these scripts')
        raise SystemExit(1)


# ---------------------------------------------------------------------------------------------------
# result files
def _h5py():
    try:
        import h5py
        return h5py
    except ImportError:
        return None


def _result_path(fname, ext):
    os.makedirs(os.path.join('results', fname[0]), exist_ok=True)
    return os.path.join('results', fname[0], f'{fname[1]}.{ext}')


def _npz_update(path, entries):
    old = dict(np.load(path, allow_pickle=False)) if os.path.exists(path) else {}
    old.update(entries)
    np.savez(path, **old)


def save_results(fname, meta=None, results=None):
    """results/<fname[0]>/<fname[1]>.h5 with /parameters attributes and /results datasets (pytvlib.py:120-139)."""
    h5py = _h5py()
    if h5py is not None:
        with h5py.File(_result_path(fname, 'h5'), 'w') as h5:
            if meta is not None:
                params = h5.create_group("parameters")
                for key, item in meta.items():
                    params.attrs[key] = item
            if results is not None:
                conv = h5.create_group("results")
                for key, item in results.items():
                    conv.create_dataset(key, dtype=np.float32, data=item)
        return _result_path(fname, 'h5')
    path = _result_path(fname, 'npz')
    if os.path.exists(path):
        os.remove(path)                                       # mode 'w' of the reference
    entries = {f'parameters@{k}': np.asarray(v) for k, v in (meta or {}).items()}
    entries.update({f'results/{k}': np.asarray(v, np.float32) for k, v in (results or {}).items()})
    _npz_update(path, entries)
    return path


def save_recon(fname, meta, tomo):
    """Append /Reconstruction/recon (float32 [Nslice, Nray, Nray]) with its size attributes (pytvlib.py:147-162).
    One bulk device-to-host copy when the engine offers get_recon_all, the reference's slice loop otherwise."""
    Nslice, Nray, Nproj = meta
    if hasattr(tomo, 'get_recon_all'):
        recon = np.asarray(tomo.get_recon_all(), np.float32)
    else:
        recon = np.zeros([Nslice, Nray, Nray], dtype=np.float32)
        for s in range(Nslice):
            recon[s, :, :] = tomo.get_recon(s)
    h5py = _h5py()
    if h5py is not None:
        with h5py.File(_result_path(fname, 'h5'), 'a') as h5:
            dset = h5.create_group("Reconstruction")
            dset.create_dataset("recon", dtype=np.float32, data=recon)
            dset.attrs["Nslice"], dset.attrs["Nray"], dset.attrs["Nproj"] = Nslice, Nray, Nproj
        return _result_path(fname, 'h5')
    path = _result_path(fname, 'npz')
    _npz_update(path, {'Reconstruction/recon': recon, 'Reconstruction@Nslice': np.asarray(Nslice),
                       'Reconstruction@Nray': np.asarray(Nray), 'Reconstruction@Nproj': np.asarray(Nproj)})
    return path

