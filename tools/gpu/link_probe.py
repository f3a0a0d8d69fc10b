"""Host-link probe: H2D rate of one step's upload (157 MB) against different amounts of concurrent D2H traffic."""
import ctypes as C, sys
sys.path.insert(0, '.')
from orb_slam_tracking_b200 import frontend as fe, capi
L = capi.lib()
E = fe.ORBextractor(1000, 1.2, 8, 20, 7, 640, 480, max_batch=512)
nin = 512 * 640 * 480
L.orbx_host_alloc.restype = C.c_void_p
hin = L.orbx_host_alloc(C.c_size_t(nin)); hout = L.orbx_host_alloc(C.c_size_t(64 << 20))
for nout in (46086144, 36 << 20, 30 << 20, 20 << 20, 4 << 20, 0):
    ms = C.c_float(0)
    capi.check(L.orbx_hostlink_probe(E._h, C.c_void_p(hin), C.c_size_t(nin), C.c_void_p(hout), C.c_size_t(nout), 8, C.byref(ms)), E._h)
    print("D2H %5.1f MB: %.3f ms per step, H2D %.1f GB/s -> ceiling %.1f k frames/s" % (nout / 1e6, ms.value, nin / ms.value / 1e6, 512 / ms.value))
