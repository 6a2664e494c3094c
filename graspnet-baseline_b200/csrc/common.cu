// common.cu -- process-wide helpers and the error-string / version entry points.
#include "common.cuh"

int b200_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return dev;
}

int b200_sm_count() {
  static std::atomic<int> cached[64];
  const int dev = b200_device();
  if (dev >= 0 && dev < 64) {
    const int c = cached[dev].load(std::memory_order_relaxed);
    if (c > 0) return c;
  }
  int sms = 0;
  if (dev < 0 || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) {
    (void)cudaGetLastError();
    return 148;  // B200
  }
  if (dev < 64) cached[dev].store(sms, std::memory_order_relaxed);
  return sms;
}

std::atomic<unsigned long long> g_b200_launches{0};

extern "C" unsigned long long b200_launch_count(void) { return g_b200_launches.load(); }

extern "C" int b200_version(void) { return 200; }

extern "C" const char *b200_error_string(int code) {
  switch (code) {
    case 0: return "success";
    case B200_EINVAL: return "b200grasp: invalid argument (null pointer or negative size)";
    case B200_ECAPACITY: return "b200grasp: problem size exceeds what this build supports";
    case B200_EWORKSPACE: return "b200grasp: workspace too small";
    case B200_ESTATIC:
      return "b200grasp: problem larger than the built-in static scratch of this no-allocation entry point (use the "
             "variant that takes a caller workspace)";
    default: break;
  }
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "b200grasp: unknown error";
}
