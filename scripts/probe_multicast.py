"""Does this box expose NVLS multicast objects (cuMulticastCreate) and fabric handles?  One JSON line; used to plan the
multicast (multimem.st / multimem.ld_reduce) variants of the push and allreduce kernels."""
import json
import torch
out = {"device": torch.cuda.get_device_name(0), "count": torch.cuda.device_count()}
try:
    from cuda import cuda
    cuda.cuInit(0)
    err, dev = cuda.cuDeviceGet(0)
    A = cuda.CUdevice_attribute
    for name in ("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED", "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_GPU_DIRECT_RDMA_SUPPORTED"):
        if hasattr(A, name):
            e, v = cuda.cuDeviceGetAttribute(getattr(A, name), dev)
            out[name.replace("CU_DEVICE_ATTRIBUTE_", "").lower()] = int(v) if int(e) == 0 else f"error {int(e)}"
        else:
            out[name] = "unknown to this cuda-python"
except Exception as e:  # noqa: BLE001
    out["error"] = repr(e)
print(json.dumps(out))
