// multi.cu — multi-GPU inside the library (SURVEY §5 "distributed communication backend", §8(b) lifecycle row, §8(e)).
//
// The reference is single-device (CudaModule::Init -> cudaSetDevice(0), CUDAModule.cpp:25; RayTracer.cu:827), so this
// layer has no counterpart there.  The BVH is replicated (every device commits the same scene), the work of an entry point
// is split, and at most ONE collective per job runs on the context's stream, right behind the kernels that feed it:
//   illumination  samples [r*S/G, (r+1)*S/G) of every triangle per rank -> ncclAllReduce(sum) of 6*nTris floats ->
//                 normalise (IlluminationEstimation.cu:125-128); or triangle ranges + broadcasts (bit-identical to 1 GPU)
//   LiDAR         contiguous ray ranges; no data-path collective (every device copies its own slice to the host), an
//                 optional gather by grouped broadcasts
//   camera        interleaved row bands (sky rows are cheap, canopy rows are not) + one group of broadcasts per buffer
// Two ways to form the communicator:
//   * rtf_create_multi(devices, n): ONE process drives n GPUs (ncclCommInitAll, one context + stream + host thread per
//     device) — what a C++ host such as UniEngine's RayTracerLayer would use;
//   * rtf_comm_init_rank(ctx, nranks, rank, id): one process per GPU (torchrun / MPI launchers), the 128-byte
//     ncclUniqueId travelling through whatever the launcher offers.
// Both end in the same per-context "sharded" entry points below.
#include <nccl.h>

#include <cstring>
#include <thread>

#include "rtf_internal.h"

namespace rtf {

static int nccl_fail(ncclResult_t r, const char* what, int line) {
  char buf[384];
  snprintf(buf, sizeof(buf), "NCCL error %d (%s) at multi.cu:%d: %s", (int)r, ncclGetErrorString(r), line, what);
  set_error(buf);
  return RTF_ERR_CUDA;
}
#define RTF_NCCL(expr)                                              \
  do {                                                              \
    ncclResult_t r_ = (expr);                                       \
    if (r_ != ncclSuccess) return nccl_fail(r_, #expr, __LINE__);   \
  } while (0)

static inline ncclComm_t comm_of(const rtf_ctx* ctx) { return (ncclComm_t)ctx->comm; }

void shard_range(uint64_t n, int nranks, int rank, uint64_t* b, uint64_t* e) {
  const uint64_t base = n / (uint64_t)nranks, rem = n % (uint64_t)nranks;
  *b = (uint64_t)rank * base + ((uint64_t)rank < rem ? (uint64_t)rank : rem);
  *e = *b + base + ((uint64_t)rank < rem ? 1u : 0u);
}

// Every rank ends up with all ranges [b_r, e_r) of `buf` (elements of `elem` bytes): one group of in-place broadcasts —
// ranges may differ in size, and no staging copy or reordering is needed, so the result is bit-identical to one GPU's.
static int gather_ranges(rtf_ctx* ctx, void* buf, uint64_t n, size_t elem) {
  if (ctx->nranks <= 1) return RTF_OK;
  RTF_NCCL(ncclGroupStart());
  for (int r = 0; r < ctx->nranks; r++) {
    uint64_t b, e;
    shard_range(n, ctx->nranks, r, &b, &e);
    if (e > b) {
      char* p = (char*)buf + b * elem;
      RTF_NCCL(ncclBroadcast(p, p, (e - b) * elem, ncclChar, r, comm_of(ctx), ctx->stream));
    }
  }
  RTF_NCCL(ncclGroupEnd());
  return RTF_OK;
}

static void phase_mark(rtf_ctx* ctx, int k) {
  if (!ctx->phase_ev[k]) cudaEventCreate(&ctx->phase_ev[k]);
  cudaEventRecord(ctx->phase_ev[k], ctx->stream);
}

// Agreement before a data collective.  Everything a rank does ahead of the collective can fail on that rank alone
// (workspace growth on a fuller device, a scene that was not committed there, a batch limit): a rank that returned early
// would leave the others blocked inside the collective for ever.  So every rank of a sharded call passes here EXACTLY ONCE,
// on every path, with its own status (>= 0 fine, < 0 its error): a 4-byte all-reduce(min) on the job's stream, read back.
// *agreed = the minimum over the ranks; the data collective runs only if that is >= 0 (the camera also skips it on 0 =
// "some rank has no acceleration structure").  Return: the rank's own error if it has one, RTF_ERR_CUDA naming the
// situation if only another rank failed, else RTF_OK.  The word lives in d_small, reserved when the communicator is formed,
// so this step itself allocates nothing.  Cost: one host round trip per sharded call (tens of microseconds).
// Not covered: a rank whose device has a sticky CUDA error cannot enqueue the agreement either; its NCCL call fails, the
// communicator is aborted locally (ncclCommAbort) and the other ranks need NCCL's own asynchronous error handling.
static int agree(rtf_ctx* ctx, int local, int* agreed) {
  *agreed = local;
  if (ctx->nranks <= 1 || !ctx->comm) return local < 0 ? local : RTF_OK;
  const std::string mine = local < 0 ? std::string(rtf_last_error()) : std::string();
  int32_t* d_word = reinterpret_cast<int32_t*>(ctx->d_small.as<char>() + 512);
  int32_t word = local;
  cudaError_t e = ctx->d_small.cap >= 1024 ? cudaMemcpyAsync(d_word, &word, 4, cudaMemcpyHostToDevice, ctx->stream)
                                           : cudaErrorInvalidValue;
  ncclResult_t r = ncclSuccess;
  if (e == cudaSuccess) r = ncclAllReduce(d_word, d_word, 1, ncclInt32, ncclMin, comm_of(ctx), ctx->stream);
  if (e == cudaSuccess && r == ncclSuccess) e = cudaMemcpyAsync(&word, d_word, 4, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && r == ncclSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess || r != ncclSuccess) {
    if (ctx->own_comm) ncclCommAbort(comm_of(ctx));  // this rank cannot take part any more: do not leave a half-entered collective
    ctx->comm = nullptr, ctx->own_comm = false;
    if (local < 0) {
      set_error(mine);
      return local;
    }
    return e != cudaSuccess ? cuda_fail(e, "agreement before the collective", __FILE__, __LINE__)
                            : nccl_fail(r, "agreement before the collective", __LINE__);
  }
  *agreed = word;
  if (local < 0) {
    set_error(mine);
    return local;
  }
  if (word < 0) {
    set_error("another rank failed before the collective (status " + std::to_string(word) + "); the job was abandoned on every rank");
    return RTF_ERR_CUDA;
  }
  return RTF_OK;
}

// The part of a sharded illumination job that runs before its collective (fallible per rank).
static int illum_sharded_compute(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push, int front,
                                 int back, int mode, float* d_out6) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  if (ctx->rng == RTF_RNG_REFERENCE_LCG) {  // as the sharded camera: the reference's sequential stream does not shard
    set_error("option rng = RTF_RNG_REFERENCE_LCG is a single-context verification mode: the sharded illumination entry points refuse it");
    return RTF_ERR_INVALID;
  }
  const uint64_t n = ctx->n_tris;
  const int faces = (front ? 1 : 0) + (back ? 1 : 0);
  int rc;
  if (mode == RTF_SHARD_SAMPLES) {
    RTF_CUDA(cudaMemsetAsync(d_out6, 0, 24 * n, ctx->stream));
    uint64_t s0, s1;
    shard_range((uint64_t)rp->samples, ctx->nranks, ctx->rank, &s0, &s1);
    if (s1 > s0)
      if ((rc = illum_mesh_device(ctx, env, rp, seed, push, front, back, 0, n, (int)s0, (int)s1, d_out6, nullptr, 0))) return rc;
  } else if (mode == RTF_SHARD_TRIANGLES) {
    uint64_t t0, t1;
    shard_range(n, ctx->nranks, ctx->rank, &t0, &t1);
    if (t1 > t0) {
      RTF_CUDA(cudaMemsetAsync(d_out6 + 6 * t0, 0, 24 * (t1 - t0), ctx->stream));
      if ((rc = illum_mesh_device(ctx, env, rp, seed, push, front, back, t0, t1, 0, rp->samples, d_out6 + 6 * t0, nullptr, 0)))
        return rc;
      if ((rc = illum_finalize_device(ctx, d_out6 + 6 * t0, t1 - t0, faces, rp->samples, nullptr))) return rc;
    }
  } else {
    set_error("mode must be RTF_SHARD_SAMPLES or RTF_SHARD_TRIANGLES");
    return RTF_ERR_INVALID;
  }
  return RTF_OK;
}

// pre_rc: status of the caller's own fallible preparation (its output buffer); non-zero skips the compute part, but the
// rank still takes part in the agreement, so that the others abandon the job instead of waiting for it.
int illum_sharded_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push, int front,
                         int back, int mode, float* d_out6, int pre_rc) {
  phase_mark(ctx, 0);
  int rc = pre_rc ? pre_rc : illum_sharded_compute(ctx, env, rp, seed, push, front, back, mode, d_out6);
  phase_mark(ctx, 1);
  int agreed;
  if ((rc = agree(ctx, rc, &agreed))) return rc;
  const uint64_t n = ctx->n_tris;
  if (mode == RTF_SHARD_SAMPLES) {
    if (ctx->nranks > 1) RTF_NCCL(ncclAllReduce(d_out6, d_out6, 6 * n, ncclFloat, ncclSum, comm_of(ctx), ctx->stream));
    phase_mark(ctx, 2);
    if ((rc = illum_finalize_device(ctx, d_out6, n, (front ? 1 : 0) + (back ? 1 : 0), rp->samples, nullptr))) return rc;
  } else {
    if ((rc = gather_ranges(ctx, d_out6, n, 24))) return rc;
    phase_mark(ctx, 2);
  }
  phase_mark(ctx, 3);
  ctx->phase_valid = true;
  return RTF_OK;
}

// Broadcast every rank's interleaved blocks of a full-size array in place (one group).
static int gather_blocks(rtf_ctx* ctx, void* buf, uint64_t total, uint32_t shift, uint64_t n_blocks, size_t elem) {
  if (ctx->nranks <= 1) return RTF_OK;
  const uint64_t B = 1ull << shift;
  RTF_NCCL(ncclGroupStart());
  for (uint64_t k = 0; k < n_blocks; k++) {
    const uint64_t b = k * B, e = (k + 1) * B < total ? (k + 1) * B : total;
    char* p = (char*)buf + b * elem;
    RTF_NCCL(ncclBroadcast(p, p, (e - b) * elem, ncclChar, (int)(k % (uint64_t)ctx->nranks), comm_of(ctx), ctx->stream));
  }
  RTF_NCCL(ncclGroupEnd());
  return RTF_OK;
}

// LiDAR shards are INTERLEAVED blocks of the ray grid (block k belongs to rank k mod nranks, ~64 blocks per rank): a
// contiguous range per rank left the ranks that scan the bare rim of a field idle while the ones over the canopy worked
// (C4 at 8 GPUs: 1.5 ms against 5.8 ms, 74 % efficiency).  Results land at the rays' own indices of the full-size arrays.
int scan_grid_sharded_device(rtf_ctx* ctx, const rtf_scan_grid* grid, int gather, float* d_t, uint32_t* d_prim, uint32_t* d_inst,
                             uint64_t* block_rays, uint64_t* n_blocks) {
  uint32_t column = 0, row = 0, shift = 0;
  int rc;
  if ((rc = rtf_scan_grid_dims(grid, &column, &row))) return rc;
  const uint64_t total = (uint64_t)column * row;
  uint64_t nb = 0, n_local = 0;
  scan_interleave_layout(total, ctx->nranks, ctx->rank, &shift, &nb, &n_local);
  if (block_rays) *block_rays = 1ull << shift;
  if (n_blocks) *n_blocks = nb;
  phase_mark(ctx, 0);
  rc = scan_grid_interleaved_device(ctx, grid, ctx->nranks, ctx->rank, true, d_t, d_prim, d_inst);
  phase_mark(ctx, 1);
  if (gather) {  // the only collective of the LiDAR path; without it a rank's failure is its own
    int agreed;
    rc = agree(ctx, rc, &agreed);
  }
  if (rc) return rc;
  if (gather) {
    if (d_t && (rc = gather_blocks(ctx, d_t, total, shift, nb, 4))) return rc;
    if (d_prim && (rc = gather_blocks(ctx, d_prim, total, shift, nb, 4))) return rc;
    if (d_inst && (rc = gather_blocks(ctx, d_inst, total, shift, nb, 4))) return rc;
  }
  phase_mark(ctx, 2);
  phase_mark(ctx, 3);
  ctx->phase_valid = true;
  return RTF_OK;
}

// Row bands of `band_rows` rows dealt round-robin: band k belongs to rank k % nranks.
static int band_rows_for(int height, int nranks) {
  if (nranks <= 1) return height > 0 ? height : 1;
  const int r = height / (16 * nranks);  // 16 bands per rank balance sky against canopy and keep batches large
  return r < 8 ? 8 : r;
}

int render_camera_sharded(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int outputs_mask) {
  if (!env || !cam || !rp) {  // no agreement possible without a camera to size the bands: argument errors are the caller's, on every rank
    set_error("rtf_render_camera_sharded: null environment, camera or ray properties");
    return RTF_ERR_INVALID;
  }
  const int h = cam->props.height, w = cam->props.width;
  const int br = band_rows_for(h, ctx->nranks);
  const int n_bands = h > 0 ? (h + br - 1) / br : 0;
  phase_mark(ctx, 0);
  int ok = render_camera_interleaved(ctx, env, cam, rp, br, ctx->nranks, ctx->rank);
  phase_mark(ctx, 1);
  if (ctx->nranks > 1) {  // the frame is gathered only if every rank rendered its bands (1); 0 on any rank: nobody gathers
    int agreed;
    const int rc = agree(ctx, ok, &agreed);
    if (rc) return rc;
    if (agreed == 0) ok = 0;
  }
  if (ok < 0) return ok;
  if (ctx->nranks > 1 && ok == 1 && w > 0) {
    float* bufs[3] = {cam->color.as<float>(), cam->normal.as<float>(), cam->albedo.as<float>()};
    for (int which = 0; which < 3; which++) {
      if (!((outputs_mask >> which) & 1)) continue;
      RTF_NCCL(ncclGroupStart());
      for (int k = 0; k < n_bands; k++) {
        const int r0 = k * br, r1 = (k + 1) * br < h ? (k + 1) * br : h;
        float* p = bufs[which] + (size_t)r0 * w * 4;
        RTF_NCCL(ncclBroadcast(p, p, (size_t)(r1 - r0) * w * 4, ncclFloat, k % ctx->nranks, comm_of(ctx), ctx->stream));
      }
      RTF_NCCL(ncclGroupEnd());
    }
    RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  phase_mark(ctx, 2);
  phase_mark(ctx, 3);
  ctx->phase_valid = true;
  return ok;
}

}  // namespace rtf

using namespace rtf;

#define CHECK_CTX(ctx)                 \
  if (!(ctx)) {                        \
    set_error("null context");         \
    return RTF_ERR_INVALID;            \
  }                                    \
  RTF_CUDA(cudaSetDevice((ctx)->device))

extern "C" {

int rtf_comm_unique_id(void* out, size_t bytes) {
  if (!out || bytes < sizeof(ncclUniqueId)) {
    set_error("rtf_comm_unique_id: the buffer must hold RTF_COMM_ID_BYTES (128) bytes");
    return RTF_ERR_INVALID;
  }
  ncclUniqueId id;
  RTF_NCCL(ncclGetUniqueId(&id));
  memcpy(out, &id, sizeof(id));
  return RTF_OK;
}
int rtf_comm_init_rank(rtf_ctx* ctx, int nranks, int rank, const void* unique_id) {
  CHECK_CTX(ctx);
  if (nranks < 1 || rank < 0 || rank >= nranks || (nranks > 1 && !unique_id)) {
    set_error("rtf_comm_init_rank: bad rank / nranks / id");
    return RTF_ERR_INVALID;
  }
  if (ctx->comm) {
    set_error("rtf_comm_init_rank: the context already has a communicator");
    return RTF_ERR_INVALID;
  }
  if (nranks == 1) {
    ctx->nranks = 1, ctx->rank = 0;
    return RTF_OK;
  }
  int rc;
  if ((rc = ctx->d_small.reserve(1024))) return rc;  // reduce values + the agreement word: nothing to allocate later
  ncclUniqueId id;
  memcpy(&id, unique_id, sizeof(id));
  ncclComm_t c = nullptr;
  RTF_NCCL(ncclCommInitRank(&c, nranks, id, rank));  // on failure the context stays a single-rank one
  ctx->comm = c, ctx->own_comm = true;
  ctx->nranks = nranks, ctx->rank = rank;
  return RTF_OK;
}
int rtf_comm_destroy(rtf_ctx* ctx) {
  if (!ctx) return RTF_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->comm && ctx->own_comm) ncclCommDestroy(comm_of(ctx));
  ctx->comm = nullptr, ctx->own_comm = false, ctx->nranks = 1, ctx->rank = 0;
  return RTF_OK;
}
int rtf_comm_rank(rtf_ctx* ctx, int* rank, int* nranks) {
  if (!ctx) return RTF_ERR_INVALID;
  if (rank) *rank = ctx->rank;
  if (nranks) *nranks = ctx->nranks;
  return RTF_OK;
}
int rtf_shard_range(uint64_t n, int nranks, int rank, uint64_t* begin, uint64_t* end) {
  if (nranks < 1 || rank < 0 || rank >= nranks || !begin || !end) return RTF_ERR_INVALID;
  shard_range(n, nranks, rank, begin, end);
  return RTF_OK;
}
// max / sum over the ranks of a few doubles (timings, ray counts): the bench's and tests' only other need for a collective
int rtf_comm_reduce_f64(rtf_ctx* ctx, double* values, int n, int op_max) {
  CHECK_CTX(ctx);
  if (!values || n < 0 || n > 64) return RTF_ERR_INVALID;
  if (ctx->nranks <= 1 || n == 0) return RTF_OK;
  int rc;
  if ((rc = ctx->d_small.reserve(64 * 8))) return rc;
  RTF_CUDA(cudaMemcpyAsync(ctx->d_small.p, values, 8 * n, cudaMemcpyHostToDevice, ctx->stream));
  RTF_NCCL(ncclAllReduce(ctx->d_small.p, ctx->d_small.p, n, ncclDouble, op_max ? ncclMax : ncclSum, comm_of(ctx), ctx->stream));
  RTF_CUDA(cudaMemcpyAsync(values, ctx->d_small.p, 8 * n, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}
int rtf_comm_barrier(rtf_ctx* ctx) {
  double one = 1.0;
  return rtf_comm_reduce_f64(ctx, &one, 1, 0);
}

int rtf_estimate_illumination_mesh_sharded_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed,
                                                  float push, int front, int back, int mode, float* dev_out6) {
  CHECK_CTX(ctx);
  if (!env || !rp || !dev_out6) return RTF_ERR_INVALID;
  return illum_sharded_device(ctx, env, rp, seed, push, front, back, mode, dev_out6, RTF_OK);
}
int rtf_estimate_illumination_mesh_sharded(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push,
                                           int front, int back, int mode, float* out6) {
  CHECK_CTX(ctx);
  if (!env || !rp || !out6) return RTF_ERR_INVALID;
  const uint64_t n = ctx->n_tris;
  int rc = ctx->d_shard_out.reserve(24 * (n ? n : 1));  // may fail on this rank alone: the others learn of it in the agreement
  if ((rc = illum_sharded_device(ctx, env, rp, seed, push, front, back, mode, ctx->d_shard_out.as<float>(), rc))) return rc;
  RTF_CUDA(cudaMemcpyAsync(out6, ctx->d_shard_out.p, 24 * n, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}
int rtf_scan_grid_trace_sharded_device(rtf_ctx* ctx, const rtf_scan_grid* grid, int gather, float* dev_t, uint32_t* dev_prim,
                                       uint32_t* dev_inst, uint64_t* block_rays, uint64_t* n_blocks) {
  CHECK_CTX(ctx);
  if (!grid || (!dev_t && !dev_prim && !dev_inst)) return RTF_ERR_INVALID;
  return scan_grid_sharded_device(ctx, grid, gather, dev_t, dev_prim, dev_inst, block_rays, n_blocks);
}
// Pure host arithmetic (no device): the block layout the sharded LiDAR entry points use for `nranks` ranks.
int rtf_scan_grid_shard_layout(const rtf_scan_grid* grid, int nranks, int rank, uint64_t* block_rays, uint64_t* n_blocks,
                               uint64_t* n_local) {
  if (!grid || nranks < 1 || rank < 0 || rank >= nranks) return RTF_ERR_INVALID;
  uint32_t column = 0, row = 0, shift = 0;
  int rc = rtf_scan_grid_dims(grid, &column, &row);
  if (rc) return rc;
  uint64_t nb = 0, nl = 0;
  scan_interleave_layout((uint64_t)column * row, nranks, rank, &shift, &nb, &nl);
  if (block_rays) *block_rays = 1ull << shift;
  if (n_blocks) *n_blocks = nb;
  if (n_local) *n_local = nl;
  return RTF_OK;
}
int rtf_render_camera_sharded(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int outputs_mask) {
  CHECK_CTX(ctx);
  return render_camera_sharded(ctx, env, cam, rp, outputs_mask);
}
// {compute, collective, finalize} milliseconds of the last sharded call on this context (device time, CUDA events)
int rtf_phase_timings(rtf_ctx* ctx, float* ms3) {
  CHECK_CTX(ctx);
  if (!ms3) return RTF_ERR_INVALID;
  ms3[0] = ms3[1] = ms3[2] = 0.0f;
  if (!ctx->phase_valid) return RTF_OK;
  RTF_CUDA(cudaEventSynchronize(ctx->phase_ev[3]));
  for (int k = 0; k < 3; k++) cudaEventElapsedTime(&ms3[k], ctx->phase_ev[k], ctx->phase_ev[k + 1]);
  return RTF_OK;
}

}  // extern "C"

// ================================ one process, n devices ===============================================================
struct rtf_multi {
  std::vector<rtf_ctx*> ctx;
};
struct rtf_multi_camera {
  rtf_multi* m = nullptr;
  std::vector<rtf_camera*> cam;  // one per device, identical properties; rank 0's buffers hold the gathered frame
};

namespace rtf {
// Runs fn(rank, ctx) on one host thread per device (NCCL collectives of one process must be entered concurrently) and
// returns the first failing status; the workers' thread-local error texts are carried over to the caller's thread.
template <class F>
static int run_all(rtf_multi* m, F fn) {
  const int n = (int)m->ctx.size();
  std::vector<int> rc(n, RTF_OK);
  std::vector<std::string> err(n);
  auto body = [&](int r) {
    rc[r] = fn(r, m->ctx[r]);
    if (rc[r] < 0) err[r] = rtf_last_error();
  };
  if (n == 1) {
    body(0);
  } else {
    std::vector<std::thread> th;
    for (int r = 0; r < n; r++) th.emplace_back(body, r);
    for (auto& t : th) t.join();
  }
  // report the rank that failed on its own account, not one that only heard of it in the agreement step
  for (int pass = 0; pass < 2; pass++)
    for (int r = 0; r < n; r++)
      if (rc[r] < 0 && (pass == 1 || err[r].rfind("another rank failed", 0) != 0)) {
        set_error("device " + std::to_string(m->ctx[r]->device) + ": " + err[r]);
        return rc[r];
      }
  return rc[0];
}
}  // namespace rtf

extern "C" {

int rtf_create_multi(const int* devices, int n, rtf_multi** out) {
  if (!out || n < 1 || n > 64 || !devices) {
    set_error("rtf_create_multi: bad arguments");
    return RTF_ERR_INVALID;
  }
  for (int a = 0; a < n; a++)
    for (int b = a + 1; b < n; b++)
      if (devices[a] == devices[b]) {
        set_error("rtf_create_multi: a device is listed twice");
        return RTF_ERR_INVALID;
      }
  rtf_multi* m = new rtf_multi();
  for (int r = 0; r < n; r++) {
    rtf_ctx* c = nullptr;
    int rc = rtf_create(devices[r], &c);
    if (rc) {
      for (rtf_ctx* x : m->ctx) rtf_destroy(x);
      delete m;
      return rc;
    }
    c->rank = r, c->nranks = n;
    m->ctx.push_back(c);
    if (n > 1 && (rc = c->d_small.reserve(1024))) {  // reduce values + the agreement word
      for (rtf_ctx* x : m->ctx) rtf_destroy(x);
      delete m;
      return rc;
    }
  }
  if (n > 1) {
    std::vector<ncclComm_t> comms(n);
    ncclResult_t r = ncclCommInitAll(comms.data(), n, devices);
    if (r != ncclSuccess) {
      for (rtf_ctx* x : m->ctx) rtf_destroy(x);
      delete m;
      return nccl_fail(r, "ncclCommInitAll", __LINE__);
    }
    for (int k = 0; k < n; k++) m->ctx[k]->comm = comms[k], m->ctx[k]->own_comm = true;
  }
  *out = m;
  return RTF_OK;
}
int rtf_multi_destroy(rtf_multi* m) {
  if (!m) return RTF_OK;
  for (rtf_ctx* c : m->ctx) rtf_comm_destroy(c), rtf_destroy(c);
  delete m;
  return RTF_OK;
}
int rtf_multi_size(rtf_multi* m, int* n) {
  if (!m || !n) return RTF_ERR_INVALID;
  *n = (int)m->ctx.size();
  return RTF_OK;
}
int rtf_multi_ctx(rtf_multi* m, int rank, rtf_ctx** ctx) {
  if (!m || !ctx || rank < 0 || rank >= (int)m->ctx.size()) return RTF_ERR_INVALID;
  *ctx = m->ctx[rank];
  return RTF_OK;
}
// ---- replicated scene sync: the same call on every device; rtf_multi_commit builds the n BVHs concurrently ----
int rtf_multi_set_geometry(rtf_multi* m, uint64_t handle, const void* vertices, size_t n_vertices, size_t stride,
                           const uint32_t* triangles, size_t n_triangles, uint64_t version) {
  if (!m) return RTF_ERR_INVALID;
  for (rtf_ctx* c : m->ctx) {
    int rc = rtf_set_geometry(c, handle, vertices, n_vertices, stride, triangles, n_triangles, version);
    if (rc) return rc;
  }
  return RTF_OK;
}
int rtf_multi_set_material(rtf_multi* m, uint64_t handle, const rtf_material* material) {
  if (!m) return RTF_ERR_INVALID;
  for (rtf_ctx* c : m->ctx) {
    int rc = rtf_set_material(c, handle, material);
    if (rc) return rc;
  }
  return RTF_OK;
}
int rtf_multi_set_instance(rtf_multi* m, uint64_t handle, uint64_t geom, uint64_t mat, const float xform[16], uint64_t comp) {
  if (!m) return RTF_ERR_INVALID;
  for (rtf_ctx* c : m->ctx) {
    int rc = rtf_set_instance(c, handle, geom, mat, xform, comp);
    if (rc) return rc;
  }
  return RTF_OK;
}
int rtf_multi_set_option(rtf_multi* m, const char* name, int64_t value) {
  if (!m) return RTF_ERR_INVALID;
  for (rtf_ctx* c : m->ctx) {
    int rc = rtf_set_option(c, name, value);
    if (rc) return rc;
  }
  return RTF_OK;
}
int rtf_multi_commit(rtf_multi* m, int morton_bits) {
  if (!m) return RTF_ERR_INVALID;
  return run_all(m, [&](int, rtf_ctx* c) { return rtf_commit(c, morton_bits); });
}

// ---- entry points ----
int rtf_multi_estimate_illumination_mesh(rtf_multi* m, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push,
                                         int front, int back, int mode, float* out6) {
  if (!m || !env || !rp || !out6) return RTF_ERR_INVALID;
  return run_all(m, [&](int r, rtf_ctx* c) -> int {
    cudaSetDevice(c->device);
    const uint64_t n = c->n_tris;
    int rc = c->d_shard_out.reserve(24 * (n ? n : 1));
    if ((rc = illum_sharded_device(c, env, rp, seed, push, front, back, mode, c->d_shard_out.as<float>(), rc))) return rc;
    if (r == 0) RTF_CUDA(cudaMemcpyAsync(out6, c->d_shard_out.p, 24 * n, cudaMemcpyDeviceToHost, c->stream));
    RTF_CUDA(cudaStreamSynchronize(c->stream));
    return RTF_OK;
  });
}
// Every device traces its interleaved blocks of the grid into dense device arrays and copies them straight to their
// places in the caller's arrays (one strided copy per array: n PCIe links in parallel, no collective — SURVEY §8e
// "preferred").
int rtf_multi_scan_grid_trace(rtf_multi* m, const rtf_scan_grid* grid, float* t_out, uint32_t* prim_out, uint32_t* inst_out) {
  if (!m || !grid) return RTF_ERR_INVALID;
  uint32_t column = 0, row = 0;
  int rc0 = rtf_scan_grid_dims(grid, &column, &row);
  if (rc0) return rc0;
  const uint64_t total = (uint64_t)column * row;
  return run_all(m, [&](int, rtf_ctx* c) -> int {
    RTF_CUDA(cudaSetDevice(c->device));
    if (!c->ready) {
      set_error("no acceleration structure (call rtf_commit)");
      return RTF_ERR_NOT_READY;
    }
    uint32_t shift = 0;
    uint64_t nb = 0, n_local = 0;
    scan_interleave_layout(total, c->nranks, c->rank, &shift, &nb, &n_local);
    if (n_local == 0) return RTF_OK;
    int rc;
    if ((rc = c->ws.io.reserve(n_local * 12))) return rc;
    float* dt = c->ws.io.as<float>();
    uint32_t* dp = (uint32_t*)(dt + n_local);
    uint32_t* di = dp + n_local;
    if ((rc = scan_grid_interleaved_device(c, grid, c->nranks, c->rank, false, t_out ? dt : nullptr, prim_out ? dp : nullptr,
                                           inst_out ? di : nullptr)))
      return rc;
    const uint64_t B = 1ull << shift;
    const uint64_t mine = ((nb - (uint64_t)c->rank) + (uint64_t)c->nranks - 1) / (uint64_t)c->nranks;  // blocks of this rank
    const bool partial_last = (nb - 1) % (uint64_t)c->nranks == (uint64_t)c->rank && nb * B != total;
    const uint64_t full = partial_last ? mine - 1 : mine;
    void* host[3] = {t_out, prim_out, inst_out};
    void* dev[3] = {dt, dp, di};
    for (int a = 0; a < 3; a++) {
      if (!host[a]) continue;
      char* h = (char*)host[a] + (uint64_t)c->rank * B * 4;
      if (full)
        RTF_CUDA(cudaMemcpy2DAsync(h, (size_t)c->nranks * B * 4, dev[a], B * 4, B * 4, full, cudaMemcpyDeviceToHost, c->stream));
      if (partial_last)
        RTF_CUDA(cudaMemcpyAsync((char*)host[a] + (nb - 1) * B * 4, (char*)dev[a] + full * B * 4, (total - (nb - 1) * B) * 4,
                                 cudaMemcpyDeviceToHost, c->stream));
    }
    RTF_CUDA(cudaStreamSynchronize(c->stream));
    return RTF_OK;
  });
}
// Host rays in, host hits out: contiguous ray ranges, each device running its own upload / trace / download pipeline.
int rtf_multi_trace_closest(rtf_multi* m, const float* rays, size_t n, void* hits) {
  if (!m || ((!rays || !hits) && n)) return RTF_ERR_INVALID;
  return run_all(m, [&](int, rtf_ctx* c) -> int {
    cudaSetDevice(c->device);
    uint64_t b, e;
    shard_range(n, c->nranks, c->rank, &b, &e);
    return rtf_trace_closest(c, rays + 8 * b, (size_t)(e - b), (char*)hits + 16 * b);
  });
}

int rtf_multi_camera_create(rtf_multi* m, rtf_multi_camera** out) {
  if (!m || !out) return RTF_ERR_INVALID;
  rtf_multi_camera* mc = new rtf_multi_camera();
  mc->m = m;
  for (rtf_ctx* c : m->ctx) {
    rtf_camera* cam = nullptr;
    int rc = rtf_camera_create(c, &cam);
    if (rc) {
      for (rtf_camera* x : mc->cam) rtf_camera_destroy(x);
      delete mc;
      return rc;
    }
    mc->cam.push_back(cam);
  }
  *out = mc;
  return RTF_OK;
}
int rtf_multi_camera_destroy(rtf_multi_camera* mc) {
  if (!mc) return RTF_OK;
  for (rtf_camera* x : mc->cam) rtf_camera_destroy(x);
  delete mc;
  return RTF_OK;
}
int rtf_multi_camera_resize(rtf_multi_camera* mc, int width, int height) {
  if (!mc) return RTF_ERR_INVALID;
  for (rtf_camera* x : mc->cam) {
    int rc = rtf_camera_resize(x, width, height);
    if (rc) return rc;
  }
  return RTF_OK;
}
int rtf_multi_camera_set(rtf_multi_camera* mc, const float position[3], const float rotation_wxyz[4]) {
  if (!mc) return RTF_ERR_INVALID;
  for (rtf_camera* x : mc->cam) {
    int rc = rtf_camera_set(x, position, rotation_wxyz);
    if (rc) return rc;
  }
  return RTF_OK;
}
// Properties of the camera on device 0; rtf_multi_render_camera copies the caller-editable fields (accumulate, fov,
// aperture, focal length, gamma, output type, max distance, modified) to the other devices before it renders.
int rtf_multi_camera_get_props(rtf_multi_camera* mc, rtf_camera_props** props) {
  if (!mc || !props || mc->cam.empty()) return RTF_ERR_INVALID;
  return rtf_camera_get_props(mc->cam[0], props);
}
int rtf_multi_render_camera(rtf_multi* m, const rtf_env* env, rtf_multi_camera* mc, const rtf_ray_props* rp, int outputs_mask) {
  if (!m || !mc || mc->m != m || !env || !rp) return RTF_ERR_INVALID;
  rtf_camera_props* p0 = nullptr;
  rtf_camera_get_props(mc->cam[0], &p0);
  for (size_t k = 1; k < mc->cam.size(); k++) {
    rtf_camera_props* pk = nullptr;
    rtf_camera_get_props(mc->cam[k], &pk);
    const int w = pk->width, h = pk->height;
    *pk = *p0;
    pk->width = w, pk->height = h;  // sizes follow rtf_multi_camera_resize (they own the device buffers)
  }
  return run_all(m, [&](int r, rtf_ctx* c) -> int {
    cudaSetDevice(c->device);
    return render_camera_sharded(c, env, mc->cam[r], rp, outputs_mask);
  });
}
int rtf_multi_camera_download(rtf_multi_camera* mc, int which, float* host) {
  if (!mc || mc->cam.empty()) return RTF_ERR_INVALID;
  return rtf_camera_download(mc->cam[0], which, host);
}

}  // extern "C"
