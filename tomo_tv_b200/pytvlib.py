"""Algorithm-name dispatch, same contract as the reference's tomofusion/pytvlib.py:5-39.

Only the dispatch helpers live here; the reference's HDF5/TIFF I/O (pytvlib.py:57-162) is out of
scope for the hot path.
"""

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
