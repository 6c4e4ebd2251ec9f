"""Probe NVLink SHARP / multicast support on this box (driver attribute + NCCL's view)."""
import ctypes, os, subprocess, sys
lib = ctypes.CDLL("libcuda.so.1")
assert lib.cuInit(0) == 0
n = ctypes.c_int(); lib.cuDeviceGetCount(ctypes.byref(n))
out = {"devices": n.value}
CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 132
CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED = 103
CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED = 128
CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED = 102
for name, attr in (("multicast", 132), ("posix_fd", 103), ("fabric", 128), ("vmm", 102)):
    vals = []
    for d in range(n.value):
        v = ctypes.c_int(-1)
        rc = lib.cuDeviceGetAttribute(ctypes.byref(v), attr, d)
        vals.append(v.value if rc == 0 else f"err{rc}")
    out[name] = vals
print(out)
# can a multicast object actually be created? (needs fabric manager with NVLS enabled)
class Prop(ctypes.Structure):
    _fields_ = [("numDevices", ctypes.c_uint), ("size", ctypes.c_size_t), ("handleTypes", ctypes.c_ulonglong), ("flags", ctypes.c_ulonglong)]
ctx = ctypes.c_void_p()
dev = ctypes.c_int(0)
lib.cuDevicePrimaryCtxRetain(ctypes.byref(ctx), 0); lib.cuCtxSetCurrent(ctx)
p = Prop(max(2, n.value) if n.value > 1 else 1, 0, 1, 0)   # handleTypes 1 = POSIX fd
gran = ctypes.c_size_t(0)
rc = lib.cuMulticastGetGranularity(ctypes.byref(gran), ctypes.byref(p), 1)  # 1 = recommended
print("cuMulticastGetGranularity rc", rc, "granularity", gran.value)
if rc == 0 and gran.value:
    p.size = gran.value
    h = ctypes.c_ulonglong(0)
    rc = lib.cuMulticastCreate(ctypes.byref(h), ctypes.byref(p))
    print("cuMulticastCreate rc", rc)
