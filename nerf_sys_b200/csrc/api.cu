// Library-level entry points: version, error string, device check.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace acn {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int num_sms_cached() {
  // per-device attribute query is cheap; cache per thread for the current device
  static thread_local int cached_dev = -1, cached_sms = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    cached_dev = dev;
    cached_sms = sms;
  }
  return cached_sms;
}

static thread_local int g_hints[ACN_HINT_COUNT] = {0};

int launch_hint(int key) { return (key >= 0 && key < ACN_HINT_COUNT) ? g_hints[key] : 0; }

}  // namespace acn

extern "C" {

int acn_launch_hint(int key, int value) {
  if (key < 0 || key >= ACN_HINT_COUNT) return -1;
  int prev = acn::g_hints[key];
  acn::g_hints[key] = value;
  return prev;
}

int acn_version(void) { return 1; }

const char* acn_last_error(void) { return acn::g_err; }

int acn_device_ok(int device) {
  int major = 0, minor = 0;
  ACN_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  ACN_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10) {
    acn::set_error("device %d is sm_%d%d; libacnerf_sm100a.so carries sm_100a code only", device, major, minor);
    return ACN_EDEVICE;
  }
  return ACN_OK;
}

int acn_num_sms(int device) {
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return -1;
  return sms;
}

}  // extern "C"
