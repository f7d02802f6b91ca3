"""Does this box expose NVLink multicast (NVLS) to user code?  Prints the device attributes the
multimem path needs."""
from cuda import cuda
err, = cuda.cuInit(0)
err, n = cuda.cuDeviceGetCount()
print("devices", n)
for d in range(n):
    err, dev = cuda.cuDeviceGet(d)
    for name in ("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED"):
        attr = getattr(cuda.CUdevice_attribute, name, None)
        if attr is None:
            print(d, name, "unknown to cuda-python")
            continue
        err, v = cuda.cuDeviceGetAttribute(attr, dev)
        print(d, name, int(v), err)
# try to create a multicast object over all devices
prop = cuda.CUmulticastObjectProp()
prop.numDevices = n
prop.size = 2 << 20
prop.handleTypes = cuda.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
prop.flags = 0
err, gran = cuda.cuMulticastGetGranularity(prop, cuda.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED)
print("granularity", err, gran)
prop.size = max(int(gran), 2 << 20) if err == cuda.CUresult.CUDA_SUCCESS else 2 << 20
err, h = cuda.cuMulticastCreate(prop)
print("cuMulticastCreate", err)
