// exchange_api.cuh — C ABI of the fused gradient exchange (exchange.cuh). Included by itmo_api.cu after train_api.cuh.
#pragma once
#include "exchange.cuh"

struct itmo_exchange {
  int device = 0, rank = 0, world = 1;
  size_t n = 0;          // floats per buffer (itmo_train_param_count)
  size_t stride = 0;     // floats between the gradient and the parameter buffer inside the window
  char* win = nullptr;   // [grads][params][flags]
  char* peer[itmo::kMaxRanks] = {};   // every rank's window as mapped into THIS process (peer[rank] == win)
  unsigned epoch = 0;
  int grid = 0;
};

namespace {
size_t xchg_window_bytes(size_t stride) { return 2 * stride * sizeof(float) + itmo::kFlagWords * sizeof(unsigned); }
float* xchg_grads(char* w) { return reinterpret_cast<float*>(w); }
float* xchg_params(char* w, size_t stride) { return reinterpret_cast<float*>(w) + stride; }
unsigned* xchg_flags(char* w, size_t stride) { return reinterpret_cast<unsigned*>(reinterpret_cast<float*>(w) + 2 * stride); }
}  // namespace

extern "C" {

int itmo_exchange_create(itmo_exchange** out, int device, void* ipc_handle_out) {
  if (!out) return fail(ITMO_ERR_INVALID, "out is null");
  *out = nullptr;
  ITMO_CUDA(cudaSetDevice(device));
  auto* x = new itmo_exchange();
  x->device = device;
  x->n = train_layout().total;
  x->stride = (x->n + 63) / 64 * 64;
  cudaError_t e = cudaMalloc((void**)&x->win, xchg_window_bytes(x->stride));
  if (e != cudaSuccess) {
    delete x;
    return fail(ITMO_ERR_NOMEM, "cudaMalloc(exchange window) failed: %s", cudaGetErrorString(e));
  }
  ITMO_CUDA(cudaMemset(x->win, 0, xchg_window_bytes(x->stride)));
  ITMO_CUDA(cudaDeviceSynchronize());
  x->peer[0] = x->win;
  if (ipc_handle_out) {
    static_assert(sizeof(cudaIpcMemHandle_t) == ITMO_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, x->win);
    if (e != cudaSuccess) {
      cudaFree(x->win);
      delete x;
      return fail(ITMO_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    memcpy(ipc_handle_out, &h, sizeof(h));
  }
  int per_sm = 0;
  ITMO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, itmo::exchange_step_kernel, 256, 0));
  x->grid = num_sms() * std::max(1, std::min(per_sm, 4));
  *out = x;
  return ITMO_OK;
}

float* itmo_exchange_grads(itmo_exchange* x) { return x ? xchg_grads(x->win) : nullptr; }
float* itmo_exchange_params(itmo_exchange* x) { return x ? xchg_params(x->win, x->stride) : nullptr; }

int itmo_exchange_open(itmo_exchange* x, int rank, int world, const void* all_handles) {
  if (!x) return fail(ITMO_ERR_INVALID, "null exchange");
  if (world < 1 || world > itmo::kMaxRanks || rank < 0 || rank >= world)
    return fail(ITMO_ERR_INVALID, "rank %d of %d: at most %d ranks of one box", rank, world, itmo::kMaxRanks);
  if (world > 1 && !all_handles) return fail(ITMO_ERR_INVALID, "handles are null");
  ITMO_CUDA(cudaSetDevice(x->device));
  x->rank = rank;
  x->world = world;
  for (int q = 0; q < itmo::kMaxRanks; ++q) x->peer[q] = nullptr;
  x->peer[rank] = x->win;
  for (int q = 0; q < world; ++q) {
    if (q == rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)all_handles + size_t(q) * ITMO_IPC_HANDLE_BYTES, sizeof(h));
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return fail(ITMO_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d) failed: %s", q, cudaGetErrorString(e));
    x->peer[q] = (char*)p;
  }
  return ITMO_OK;
}

int itmo_exchange_step(itmo_exchange* x, itmo_trainer* t, int optimizer, float lr, float beta1_or_rho, float beta2,
                       float eps, float grad_scale, void* stream) {
  if (!x || !t) return fail(ITMO_ERR_INVALID, "null argument");
  if (t->params != xchg_params(x->win, x->stride) || t->grads != xchg_grads(x->win))
    return fail(ITMO_ERR_STATE, "the trainer's flat buffers are not this exchange's window (itmo_exchange_params / _grads)");
  if (optimizer != ITMO_OPT_RMSPROP && optimizer != ITMO_OPT_ADAM) return fail(ITMO_ERR_INVALID, "unknown optimizer %d", optimizer);
  for (int q = 0; q < x->world; ++q)
    if (!x->peer[q]) return fail(ITMO_ERR_STATE, "itmo_exchange_step before itmo_exchange_open");
  ITMO_CUDA(cudaSetDevice(x->device));
  t->opt_steps++;
  itmo::ExchangeArgs a;
  memset(&a, 0, sizeof(a));
  for (int q = 0; q < x->world; ++q) {
    a.grads[q] = xchg_grads(x->peer[q]);
    a.params[q] = xchg_params(x->peer[q], x->stride);
    a.flags[q] = xchg_flags(x->peer[q], x->stride);
  }
  a.state0 = t->opt_state;
  a.state1 = t->opt_state + x->n;
  a.n = x->n;
  const unsigned long long per = ((x->n + x->world - 1) / x->world + 3) / 4 * 4;
  a.begin = std::min<unsigned long long>(x->n, per * x->rank);
  a.end = std::min<unsigned long long>(x->n, a.begin + per);
  a.rank = x->rank;
  a.world = x->world;
  a.epoch = x->epoch++;
  a.adam = optimizer == ITMO_OPT_ADAM;
  a.lr = lr;
  if (a.adam) {
    const double b1t = pow(double(beta1_or_rho), double(t->opt_steps)), b2t = pow(double(beta2), double(t->opt_steps));
    a.lr = float(double(lr) * sqrt(1.0 - b2t) / (1.0 - b1t));
  }
  a.b1 = beta1_or_rho;
  a.b2 = beta2;
  a.eps = eps;
  a.gscale = grad_scale;
  void* args[] = {&a};
  ITMO_CUDA(cudaLaunchCooperativeKernel((const void*)itmo::exchange_step_kernel, dim3(x->grid), dim3(256), args, 0,
                                        (cudaStream_t)stream));
  t->launches++;
  return itmo_trainer_sync_weights(t, stream);
}

void itmo_exchange_destroy(itmo_exchange* x) {
  if (!x) return;
  cudaSetDevice(x->device);
  cudaDeviceSynchronize();
  for (int q = 0; q < itmo::kMaxRanks; ++q)
    if (x->peer[q] && x->peer[q] != x->win) cudaIpcCloseMemHandle(x->peer[q]);
  cudaFree(x->win);
  delete x;
}

}  // extern "C"
