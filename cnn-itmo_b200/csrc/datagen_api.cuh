// datagen_api.cuh — C ABI over datagen_ops.cuh (include/itmo_b200.h, "training-data synthesis and augmentation").
// Included at the end of itmo_api.cu (same translation unit: fail(), ITMO_CUDA, num_sms()).
#pragma once
#include "datagen_ops.cuh"

namespace {

int dg_grid(size_t items) { return int(std::min<size_t>((items + 255) / 256, size_t(num_sms()) * 8)); }

// G[img] for n images on `s`; returns device pointers allocated stream-ordered (freed by the caller with cudaFreeAsync)
int dg_geomean(const float* hdr, int n, int h, int w, double** G_out, double** partial_out, cudaStream_t s) {
  double *partial = nullptr, *G = nullptr;
  ITMO_CUDA(cudaMallocAsync((void**)&partial, sizeof(double) * size_t(n) * itmo::kDgBlocksPerImage, s));
  cudaError_t e = cudaMallocAsync((void**)&G, sizeof(double) * n, s);
  if (e != cudaSuccess) {
    cudaFreeAsync(partial, s);
    return fail(ITMO_ERR_NOMEM, "cudaMallocAsync failed: %s", cudaGetErrorString(e));
  }
  itmo::loglum_partial_kernel<<<dim3(itmo::kDgBlocksPerImage, n), 256, 0, s>>>(hdr, h, w, partial);
  itmo::geomean_finalize_kernel<<<(n + 127) / 128, 128, 0, s>>>(partial, itmo::kDgBlocksPerImage, double(h) * w, n, G);
  *G_out = G;
  *partial_out = partial;
  ITMO_CUDA(cudaGetLastError());
  return ITMO_OK;
}

int dg_tonemap(const float* hdr, int n, int h, int w, const itmo_camera_params* cams_host, int mode, double key, void* out,
               int out_dtype, cudaStream_t s) {
  if (!hdr || !out) return fail(ITMO_ERR_INVALID, "null argument");
  if (n <= 0 || h <= 0 || w <= 0) return fail(ITMO_ERR_INVALID, "bad shape");
  if (out_dtype != ITMO_DT_U8 && out_dtype != ITMO_DT_F32) return fail(ITMO_ERR_INVALID, "bad out_dtype");
  double *G = nullptr, *partial = nullptr;
  int rc = dg_geomean(hdr, n, h, w, &G, &partial, s);
  if (rc) return rc;
  itmo::CameraParams* cam = nullptr;
  if (mode == 0) {
    static_assert(sizeof(itmo::CameraParams) == sizeof(itmo_camera_params), "ABI struct mismatch");
    cudaError_t e = cudaMallocAsync((void**)&cam, sizeof(itmo::CameraParams) * n, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(cam, cams_host, sizeof(itmo::CameraParams) * n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);   // cams_host may be a temporary of the caller
    if (e != cudaSuccess) {
      cudaFreeAsync(G, s);
      cudaFreeAsync(partial, s);
      if (cam) cudaFreeAsync(cam, s);
      return fail(ITMO_ERR_CUDA, "camera parameter upload failed: %s", cudaGetErrorString(e));
    }
  }
  const dim3 grid(std::max(1, std::min(dg_grid(size_t(h) * w), std::max(1, num_sms() * 8 / n))), n);
  if (out_dtype == ITMO_DT_U8)
    itmo::tonemap_kernel<uint8_t><<<grid, 256, 0, s>>>(hdr, (uint8_t*)out, h, w, G, cam, mode, key);
  else
    itmo::tonemap_kernel<float><<<grid, 256, 0, s>>>(hdr, (float*)out, h, w, G, cam, mode, key);
  cudaError_t e = cudaGetLastError();
  cudaFreeAsync(G, s);
  cudaFreeAsync(partial, s);
  if (cam) cudaFreeAsync(cam, s);
  if (e != cudaSuccess) return fail(ITMO_ERR_CUDA, "tonemap launch failed: %s", cudaGetErrorString(e));
  return ITMO_OK;
}

}  // namespace

extern "C" {

int itmo_synth_hdr_field(float* hdr_dev, int n, int h, int w, unsigned seed, void* stream) {
  if (!hdr_dev) return fail(ITMO_ERR_INVALID, "null argument");
  if (n <= 0 || h <= 0 || w <= 0) return fail(ITMO_ERR_INVALID, "bad shape");
  if (size_t(n) * h * w * 3 >= (size_t(1) << 32)) return fail(ITMO_ERR_INVALID, "field too large for the 32-bit counter");
  itmo::synth_hdr_kernel<<<dg_grid(size_t(n) * h * w * 3), 256, 0, (cudaStream_t)stream>>>(hdr_dev, n, h, w, seed);
  ITMO_CUDA(cudaGetLastError());
  return ITMO_OK;
}

int itmo_virtual_camera(const float* hdr_dev, int n, int h, int w, const itmo_camera_params* cams_host, void* sdr_dev,
                        int out_dtype, void* stream) {
  if (!cams_host) return fail(ITMO_ERR_INVALID, "null argument");
  return dg_tonemap(hdr_dev, n, h, w, cams_host, 0, 0.18, sdr_dev, out_dtype, (cudaStream_t)stream);
}

int itmo_reinhard_target(const float* hdr_dev, int n, int h, int w, double key, void* out_dev, int out_dtype,
                         void* stream) {
  return dg_tonemap(hdr_dev, n, h, w, nullptr, 1, key, out_dev, out_dtype, (cudaStream_t)stream);
}

int itmo_augment(const void* in_dev, int in_dtype, int n, int h, int w, const itmo_affine_params* params_host,
                 float* out_dev, float scale, void* stream) {
  if (!in_dev || !params_host || !out_dev) return fail(ITMO_ERR_INVALID, "null argument");
  if (n <= 0 || h <= 1 || w <= 1) return fail(ITMO_ERR_INVALID, "bad shape");
  if (in_dtype != ITMO_DT_U8 && in_dtype != ITMO_DT_F32) return fail(ITMO_ERR_INVALID, "bad in_dtype");
  static_assert(sizeof(itmo::AffineParams) == sizeof(itmo_affine_params), "ABI struct mismatch");
  cudaStream_t s = (cudaStream_t)stream;
  itmo::AffineParams* P = nullptr;
  ITMO_CUDA(cudaMallocAsync((void**)&P, sizeof(itmo::AffineParams) * n, s));
  cudaError_t e = cudaMemcpyAsync(P, params_host, sizeof(itmo::AffineParams) * n, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) {
    cudaFreeAsync(P, s);
    return fail(ITMO_ERR_CUDA, "augmentation parameter upload failed: %s", cudaGetErrorString(e));
  }
  const dim3 grid(std::max(1, std::min(dg_grid(size_t(h) * w), std::max(1, num_sms() * 8 / n))), n);
  if (in_dtype == ITMO_DT_U8)
    itmo::augment_kernel<uint8_t><<<grid, 256, 0, s>>>((const uint8_t*)in_dev, out_dev, h, w, P, scale);
  else
    itmo::augment_kernel<float><<<grid, 256, 0, s>>>((const float*)in_dev, out_dev, h, w, P, scale);
  e = cudaGetLastError();
  cudaFreeAsync(P, s);
  if (e != cudaSuccess) return fail(ITMO_ERR_CUDA, "augment launch failed: %s", cudaGetErrorString(e));
  return ITMO_OK;
}

}  // extern "C"
