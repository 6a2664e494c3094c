// common.cu -- process-wide helpers and the error-string / version entry points.
#include "common.cuh"

int b200_sm_count() {
  static int cached = 0;
  if (cached > 0) return cached;
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) {
    (void)cudaGetLastError();
    return 148;  // B200
  }
  cached = sms;
  return sms;
}

unsigned long long g_b200_launches = 0;

extern "C" unsigned long long b200_launch_count(void) { return g_b200_launches; }

extern "C" int b200_version(void) { return 100; }

extern "C" const char *b200_error_string(int code) {
  switch (code) {
    case 0: return "success";
    case B200_EINVAL: return "b200grasp: invalid argument (null pointer or negative size)";
    case B200_ECAPACITY: return "b200grasp: problem size exceeds what this build supports";
    case B200_EWORKSPACE: return "b200grasp: workspace too small";
    default: break;
  }
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "b200grasp: unknown error";
}
