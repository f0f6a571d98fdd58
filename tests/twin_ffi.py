"""ctypes binding of oracle/libslamtwin.so (test infrastructure only)."""
import ctypes as C
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int)
i64p = C.POINTER(C.c_int64)


def load_twin():
    lib = C.CDLL(str(ROOT / "oracle" / "libslamtwin.so"))
    lib.twin_mpi_to_pi.restype = C.c_double
    lib.twin_mpi_to_pi.argtypes = [C.c_double]
    lib.twin_predict.restype = None
    lib.twin_predict.argtypes = [dp, dp, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, dp, C.c_double]
    lib.twin_compute_association.restype = None
    lib.twin_compute_association.argtypes = [dp, dp, C.c_int, C.c_int, dp, dp, C.c_int64, dp, dp]
    lib.twin_associate.restype = None
    lib.twin_associate.argtypes = [dp, dp, C.c_int, C.c_int, dp, C.c_int, dp, C.c_double, C.c_double, dp,
                                   i64p, ip, dp, ip, dp]
    lib.twin_update.restype = C.c_int
    lib.twin_update.argtypes = [dp, dp, C.c_int, C.c_int, dp, C.c_int, dp, i64p]
    lib.twin_add_features.restype = C.c_int
    lib.twin_add_features.argtypes = [dp, dp, C.c_int, C.c_int, dp, C.c_int, dp]
    lib.twin_normal.restype = C.c_double
    lib.twin_normal.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
    lib.twin_sim.restype = C.c_int
    lib.twin_sim.argtypes = [dp, C.c_int, dp, C.c_int, dp, dp, C.c_int, C.c_uint64, C.c_uint32, C.c_int,
                             C.c_int, dp, dp, dp, dp, ip, dp]
    lib.twin_ensemble.restype = C.c_int
    lib.twin_ensemble.argtypes = [dp, C.c_int, dp, C.c_int, C.c_uint64, C.c_uint32, C.c_int, C.c_int,
                                  C.c_int, dp, dp, ip, C.c_int]
    return lib


def d(a):
    return a.ctypes.data_as(dp)


class Twin:
    """numpy-level wrappers with the oracle's (x, P, ...) signatures; P is n x n symmetric."""

    def __init__(self, lib=None):
        self.lib = lib or load_twin()

    def predict(self, x, P, g, v, w, Q, dt):
        x = np.array(x, dtype=float)
        Pf = np.asfortranarray(np.array(P, dtype=float))
        n = len(x)
        q = np.ascontiguousarray(np.asarray(Q, dtype=float).T).reshape(-1)
        self.lib.twin_predict(d(x), d(Pf), n, n, g, v, w, d(q), dt)
        return x, np.ascontiguousarray(Pf)

    def compute_association(self, x, P, z, R, idf):
        x = np.ascontiguousarray(x, dtype=float)
        Pf = np.asfortranarray(np.asarray(P, dtype=float))
        n = len(x)
        z = np.ascontiguousarray(z, dtype=float)
        r = np.ascontiguousarray(np.asarray(R, dtype=float).T).reshape(-1)
        nis, nd = C.c_double(), C.c_double()
        self.lib.twin_compute_association(d(x), d(Pf), n, n, d(z), d(r), int(idf), C.byref(nis), C.byref(nd))
        return nis.value, nd.value

    def associate(self, x, P, z, R, gate1, gate2):
        x = np.ascontiguousarray(x, dtype=float)
        Pf = np.asfortranarray(np.asarray(P, dtype=float))
        n = len(x)
        z = np.asarray(z, dtype=float).reshape(2, -1)
        nz = z.shape[1]
        zc = np.ascontiguousarray(z.T).reshape(-1)
        r = np.ascontiguousarray(np.asarray(R, dtype=float).T).reshape(-1)
        zf = np.zeros(2 * max(nz, 1))
        zn = np.zeros(2 * max(nz, 1))
        idf = np.zeros(max(nz, 1), dtype=np.int64)
        m, k, mg = C.c_int(), C.c_int(), C.c_double()
        self.lib.twin_associate(d(x), d(Pf), n, n, d(zc), nz, d(r), gate1, gate2, d(zf),
                                idf.ctypes.data_as(i64p), C.byref(m), d(zn), C.byref(k), C.byref(mg))
        m, k = m.value, k.value
        self.last_margin = mg.value
        return zf[:2 * m].reshape(m, 2).T.copy(), idf[:m].copy(), zn[:2 * k].reshape(k, 2).T.copy()

    def update(self, x, P, zf, R, idf):
        x = np.array(x, dtype=float)
        Pf = np.asfortranarray(np.array(P, dtype=float))
        n = len(x)
        zf = np.asarray(zf, dtype=float).reshape(2, -1)
        m = zf.shape[1]
        zc = np.ascontiguousarray(zf.T).reshape(-1)
        ids = np.ascontiguousarray(np.asarray(idf, dtype=np.int64).reshape(-1))
        r = np.ascontiguousarray(np.asarray(R, dtype=float).T).reshape(-1)
        rc = self.lib.twin_update(d(x), d(Pf), n, n, d(zc), m, d(r), ids.ctypes.data_as(i64p))
        if rc != 0:
            raise np.linalg.LinAlgError("twin_update: S not PD")
        return x, np.ascontiguousarray(Pf)

    def add_features(self, x, P, zn, R):
        zn = np.asarray(zn, dtype=float).reshape(2, -1)
        k = zn.shape[1]
        n = len(x)
        n2 = n + 2 * k
        xx = np.zeros(n2)
        xx[:n] = x
        Pf = np.zeros((n2, n2), order="F")
        Pf[:n, :n] = P
        zc = np.ascontiguousarray(zn.T).reshape(-1)
        r = np.ascontiguousarray(np.asarray(R, dtype=float).T).reshape(-1)
        nn = self.lib.twin_add_features(d(xx), d(Pf), n, n2, d(zc), k, d(r))
        assert nn == n2
        return xx, np.ascontiguousarray(Pf)
