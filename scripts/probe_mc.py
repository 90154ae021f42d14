import os, sys
try:
    from cuda.bindings import driver as cu
except Exception:
    from cuda import cuda as cu
def ck(r):
    if isinstance(r, tuple):
        err = r[0]; rest = r[1:]
    else:
        err, rest = r, ()
    if err != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(str(err))
    return rest[0] if len(rest) == 1 else rest
ck(cu.cuInit(0))
n = ck(cu.cuDeviceGetCount())
print("devices", n)
for d in range(n):
    dev = ck(cu.cuDeviceGet(d))
    mc = ck(cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev))
    vmm = ck(cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev))
    fdh = ck(cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev))
    fab = ck(cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev))
    print(d, "multicast", mc, "vmm", vmm, "posix_fd", fdh, "fabric", fab)
prop = cu.CUmulticastObjectProp()
prop.numDevices = n
prop.size = 2 << 20
prop.handleTypes = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
prop.flags = 0
try:
    g = ck(cu.cuMulticastGetGranularity(prop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED))
    print("mc granularity recommended", g)
    import torch
    torch.cuda.init(); torch.zeros(1, device="cuda")
    h = ck(cu.cuMulticastCreate(prop))
    print("cuMulticastCreate ok", h)
except Exception as e:
    print("multicast create failed:", e)
import ctypes
print("pidfd_open avail", hasattr(os, "pidfd_open"))
