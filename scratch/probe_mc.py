import sys
try:
    from cuda import cuda
except Exception:
    from cuda.bindings import driver as cuda
err, = cuda.cuInit(0)
err, n = cuda.cuDeviceGetCount()
print("devices", n)
for i in range(n):
    err, dev = cuda.cuDeviceGet(i)
    for name in ["CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED", "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED"]:
        a = getattr(cuda.CUdevice_attribute, name, None)
        if a is None:
            print(i, name, "n/a"); continue
        err, v = cuda.cuDeviceGetAttribute(a, dev)
        print(i, name, v, err)
print(open("/proc/sys/kernel/yama/ptrace_scope").read().strip() if __import__("os").path.exists("/proc/sys/kernel/yama/ptrace_scope") else "no yama")
