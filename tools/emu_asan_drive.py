"""Driver of tools/emu_asan.sh: calls the emulated NCC / SSIM kernels (tests/native/*_emulated.cpp built with
-fsanitize=address) on exactly-sized malloc'd buffers, so any out-of-bounds global or shared-memory access aborts."""
import sys
import ctypes as C, numpy as np, math
# numpy buffers allocated by malloc -> ASan sees their bounds
def arr(n, fill=None, dtype=np.float32):
    a = np.empty(n, dtype=dtype)
    if fill is not None: a[:] = fill
    return a
p = lambda a: C.c_void_p(a.ctypes.data)
L = C.CDLL(sys.argv[1] + "/ssim_emu_asan.so")
g = np.array([math.exp(-((x-5)**2)/(2*1.5**2)) for x in range(11)], dtype=np.float32); g /= g.sum()
rng = np.random.default_rng(0)
for (planes,H,W,valid) in ((3,40,50,1),(3,40,50,0),(2,33,64,1),(1,75,70,0),(2,12,11,1),(1,5,7,0)):
    n = planes*H*W
    i1 = rng.random(n, dtype=np.float32); i2 = rng.random(n, dtype=np.float32)
    dm = [arr(n) for _ in range(3)]
    rh, rw = (H-10, W-10) if valid else (H, W)
    nt = ((rh+31)//32)*((rw+31)//32)*planes
    partials = arr(nt); mean = arr(1); gm = arr(1, 0.5); out = arr(n)
    L.ggs_test_ssim_forward(planes,H,W,valid,p(g),p(i1),p(i2),p(dm[0]),p(dm[1]),p(dm[2]),p(partials),p(mean))
    L.ggs_test_ssim_backward(planes,H,W,valid,p(g),p(i1),p(i2),p(dm[0]),p(dm[1]),p(dm[2]),p(gm),p(out))
    print("ssim ok", planes,H,W,valid, float(mean[0]))
L2 = C.CDLL(sys.argv[1] + "/ncc_emu_asan.so")
H,W = 60, 80
for n in (1, 129, 300):
    ir = rng.random(H*W, dtype=np.float32); im = rng.random(H*W, dtype=np.float32)
    uv = np.stack([rng.integers(0,W,n), rng.integers(0,H,n)],1).astype(np.int32).copy()
    d = (2+rng.random(n)).astype(np.float32); nr = np.tile(np.array([0,0,-1],np.float32),(n,1)).copy()
    R = np.eye(3,dtype=np.float32).ravel().copy(); T = np.array([0.03,-0.02,0.01],np.float32)
    intr = np.array([150,150,W/2,H/2,150,150,W/2,H/2],np.float32)
    o = [arr(n), arr(n), arr(3*n), arr(n, dtype=np.uint8)]
    L2.ggs_test_ncc_kernel(C.c_longlong(n), p(d), p(nr), p(uv), p(R), p(T), p(ir), p(im), p(intr), H,W,H,W, p(o[0]),p(o[1]),p(o[2]),p(o[3]))
    print("ncc ok", n, int(o[3].sum()))
