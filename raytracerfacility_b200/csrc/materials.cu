// materials.cu — the table-shaped material inputs behind the C ABI: images in linear memory (textures, cube-map
// faces) and compressed-BTF databases, plus their test hooks.
//
// Reference: RayTracedMaterial's GL texture ids bound through CUDA-GL interop (RT.hpp:254-263, RT.cu:2233-2256), the
// cube-map faces of EnvironmentProperties (RT.hpp:112-113, RT.cu:270-347), CompressedBTF::ImportFromFolder /
// UploadDeviceData (src/RayTracerFacility/MLVQ/CompressedBTF.cpp:74-390, :420-520).  No GL here: images and tables are
// uploaded from host arrays into plain device memory.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>

#include "rtf_internal.h"
#include "rtf_shade.cuh"

namespace rtf {

#define CHECK_CTX(ctx)                 \
  if (!(ctx)) {                        \
    set_error("null context");         \
    return RTF_ERR_INVALID;            \
  }                                    \
  RTF_CUDA(cudaSetDevice((ctx)->device))

// ---- BTF tables -------------------------------------------------------------------------------------------------------
void BTFTables::release() {
  DevBuf* bufs[] = {&pdf6_slices, &pdf6_scales, &pdf1d_basis, &colors,       &index_ab,     &pdf2d_colors, &pdf2d_slices,
                    &pdf2d_scales, &pdf2d_indices, &pdf3d_scales, &pdf3d_slices, &pdf4d_scales, &pdf4d_slices, &beta_angles};
  for (DevBuf* b : bufs) b->release();
}
DevBTF BTFTables::dev() const {
  DevBTF d;
  d.use_cos_beta = desc.use_cos_beta, d.num_beta = desc.num_beta, d.num_alpha = desc.num_alpha;
  d.num_theta = desc.num_theta, d.num_phi = desc.num_phi, d.code_btf_flag = desc.code_btf_flag, d.hdr = desc.hdr;
  d.hdr_value = desc.hdr_value, d.multiplier = desc.multiplier, d.tex_coord_multiplier = desc.tex_coord_multiplier;
  d.gamma = desc.gamma;
  d.rows_offset = desc.rows_offset, d.cols_offset = desc.cols_offset, d.pdf6_rows = desc.pdf6_rows, d.pdf6_cols = desc.pdf6_cols;
  d.n_channels = desc.n_channels, d.pdf2d_slice_len = desc.pdf2d_slice_len;
  d.pdf6_slices = pdf6_slices.as<int>(), d.pdf6_scales = pdf6_scales.as<float>(), d.pdf1d_basis = pdf1d_basis.as<float>();
  d.colors = colors.as<float>(), d.index_ab = index_ab.as<int>(), d.pdf2d_colors = pdf2d_colors.as<int>();
  d.pdf2d_slices = pdf2d_slices.as<int>(), d.pdf2d_scales = pdf2d_scales.as<float>(), d.pdf2d_indices = pdf2d_indices.as<int>();
  d.pdf3d_scales = pdf3d_scales.as<float>(), d.pdf3d_slices = pdf3d_slices.as<int>(), d.pdf4d_scales = pdf4d_scales.as<float>();
  d.pdf4d_slices = pdf4d_slices.as<int>(), d.beta_angles = beta_angles.as<float>();
  return d;
}

void release_material_tables(rtf_ctx* ctx) {
  for (auto& t : ctx->textures) t.second.texels.release();
  for (auto& b : ctx->btfs) b.second.release();
  ctx->textures.clear(), ctx->btfs.clear();
}

static DevTexture dev_texture(const Texture& t) {
  DevTexture d;
  d.texels = t.texels.as<float4>(), d.w = t.w, d.h = t.h;
  return d;
}

// materials -> dense DevMaterial table with resolved texture / BTF pointers (called by rtf_commit)
int resolve_materials(rtf_ctx* ctx, std::map<uint64_t, uint32_t>& mat_index, std::vector<DevMaterial>& mats,
                      std::vector<DevBTF>& btfs) {
  std::map<uint64_t, int> btf_index;
  for (auto& kv : ctx->mats) {
    DevMaterial dm;
    dm.m = kv.second;
    dm.btf = -1;
    dm.rich = dm.m.material_type != RTF_MAT_DEFAULT;
    auto lk = ctx->mat_links.find(kv.first);
    for (int k = 0; k < 4; k++) {
      const uint64_t th = lk == ctx->mat_links.end() ? 0 : lk->second.tex[k];
      if (th == 0) continue;
      auto t = ctx->textures.find(th);
      if (t == ctx->textures.end()) {
        set_error("material references an unknown texture");
        return RTF_ERR_NOT_FOUND;
      }
      dm.tex[k] = dev_texture(t->second);
      dm.rich = 1;
    }
    if (dm.m.material_type == RTF_MAT_COMPRESSED_BTF) {
      const uint64_t bh = lk == ctx->mat_links.end() ? 0 : lk->second.btf;
      auto b = ctx->btfs.find(bh);
      if (bh == 0 || b == ctx->btfs.end()) {
        set_error("CompressedBTF material without a BTF (rtf_set_material_btf)");
        return RTF_ERR_NOT_FOUND;
      }
      auto bi = btf_index.find(bh);
      if (bi == btf_index.end()) {
        bi = btf_index.emplace(bh, (int)btfs.size()).first;
        btfs.push_back(b->second.dev());
      }
      dm.btf = bi->second;
    }
    mat_index[kv.first] = (uint32_t)mats.size();
    mats.push_back(dm);
  }
  return RTF_OK;
}

int resolve_environment_map(rtf_ctx* ctx, const DevTexture** dev_faces) {
  *dev_faces = nullptr;
  if (!ctx->has_env_map) return RTF_OK;
  DevTexture faces[6];
  for (int k = 0; k < 6; k++) {
    auto t = ctx->textures.find(ctx->env_map[k]);
    if (t == ctx->textures.end()) {
      set_error("environment map references an unknown texture");
      return RTF_ERR_NOT_FOUND;
    }
    faces[k] = dev_texture(t->second);
  }
  int rc;
  if ((rc = ctx->d_env_faces.reserve(sizeof(faces)))) return rc;
  RTF_CUDA(cudaMemcpyAsync(ctx->d_env_faces.p, faces, sizeof(faces), cudaMemcpyHostToDevice, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));  // `faces` is a stack array
  *dev_faces = ctx->d_env_faces.as<DevTexture>();
  return RTF_OK;
}

// ---- test-hook kernels --------------------------------------------------------------------------------------------------
__global__ void k_sample_texture(DevTexture tex, uint32_t n, const float* __restrict__ uv, float4* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = tex2d(tex, uv[2 * i], uv[2 * i + 1]);
}
__global__ void k_sample_environment(DevScene sc, rtf_env env, uint32_t n, const float* __restrict__ pd, float* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* q = pd + 6 * (size_t)i;
  const float3 c = env_light(f3(q[0], q[1], q[2]), f3(q[3], q[4], q[5]), env, sc.env_faces);
  out[3 * (size_t)i] = c.x, out[3 * (size_t)i + 1] = c.y, out[3 * (size_t)i + 2] = c.z;
}
// in14 per query: tex_coord 2, view 3, illum 3, normal 3, tangent 3
__global__ void k_btf_eval(DevBTF btf, uint32_t n, const float* __restrict__ in14, float* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* q = in14 + 14 * (size_t)i;
  const float3 c = btf_get_value<true>(btf, q[0], q[1], f3(q[2], q[3], q[4]), f3(q[5], q[6], q[7]), f3(q[8], q[9], q[10]),
                                 f3(q[11], q[12], q[13]));
  out[3 * (size_t)i] = c.x, out[3 * (size_t)i + 1] = c.y, out[3 * (size_t)i + 2] = c.z;
}

}  // namespace rtf

using namespace rtf;

static int upload(DevBuf& buf, const void* host, size_t bytes, cudaStream_t st) {
  int rc;
  if ((rc = buf.reserve(bytes))) return rc;
  if (bytes) RTF_CUDA(cudaMemcpyAsync(buf.p, host, bytes, cudaMemcpyHostToDevice, st));
  return RTF_OK;
}
static bool indices_in_range(const int32_t* v, size_t n, size_t stride, int lo, int hi_excl) {
  for (size_t i = 0; i < n; i += stride)
    if (v[i] < lo || v[i] >= hi_excl) return false;
  return true;
}

// ---- CompressedBTF::ImportFromFolder (CompressedBTF.cpp:74-390) -----------------------------------------------------------
// "rows cols min max" header + row-major values (ParseFloatData / ParseIntData, CompressedBTF.cpp:7-50)
template <class T>
static bool parse_table(const std::string& path, int& rows, int& cols, std::vector<T>& data) {
  FILE* fp = fopen(path.c_str(), "r");
  if (!fp) {
    set_error("cannot open " + path);
    return false;
  }
  double mn, mx;
  bool ok = fscanf(fp, "%d %d %lf %lf", &rows, &cols, &mn, &mx) == 4 && rows >= 0 && cols >= 0 &&
            (size_t)rows * (size_t)cols < ((size_t)1 << 31);
  if (ok) {
    data.resize((size_t)rows * cols);
    for (size_t i = 0; i < data.size() && ok; i++) {
      double v;
      ok = fscanf(fp, "%lf", &v) == 1;
      data[i] = (T)v;
    }
  }
  fclose(fp);
  if (!ok) set_error("corrupt table file " + path);
  return ok;
}

extern "C" {

int rtf_set_texture(rtf_ctx* ctx, uint64_t handle, int width, int height, const float* rgba) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  if (handle == 0 || width < 1 || height < 1 || !rgba || (size_t)width * height > ((size_t)1 << 28)) {
    set_error("rtf_set_texture: bad handle, size or pixels");
    return RTF_ERR_INVALID;
  }
  // An existing handle may back the committed material / cube-map tables (old pointer and size): uploading over it, or
  // re-allocating it, would leave them dangling until the next commit.  Same guard as rtf_remove_texture.
  if (ctx->textures.count(handle)) {
    RTF_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->ready = false;
  }
  Texture& t = ctx->textures[handle];
  t.w = width, t.h = height;
  int rc;
  if ((rc = upload(t.texels, rgba, 16 * (size_t)width * height, ctx->stream))) return rc;
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));  // the caller's array may die at return
  return RTF_OK;
}
int rtf_remove_texture(rtf_ctx* ctx, uint64_t handle) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  auto it = ctx->textures.find(handle);
  if (it == ctx->textures.end()) return RTF_ERR_NOT_FOUND;
  cudaStreamSynchronize(ctx->stream);
  it->second.texels.release();
  ctx->textures.erase(it);
  ctx->ready = false;  // committed materials may point at it: a new rtf_commit is needed
  return RTF_OK;
}
int rtf_set_material_textures(rtf_ctx* ctx, uint64_t material, uint64_t albedo, uint64_t normal, uint64_t metallic,
                              uint64_t roughness) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  MaterialLinks& l = ctx->mat_links[material];
  l.tex[MAT_TEX_ALBEDO] = albedo, l.tex[MAT_TEX_NORMAL] = normal, l.tex[MAT_TEX_METALLIC] = metallic;
  l.tex[MAT_TEX_ROUGHNESS] = roughness;
  return RTF_OK;
}
int rtf_set_environment_map(rtf_ctx* ctx, const uint64_t faces[6]) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  ctx->has_env_map = faces != nullptr;
  for (int k = 0; k < 6; k++) ctx->env_map[k] = faces ? faces[k] : 0;
  return RTF_OK;
}
int rtf_sample_texture(rtf_ctx* ctx, uint64_t handle, size_t n, const float* uv, float* rgba_out) {
  CHECK_CTX(ctx);
  auto it = ctx->textures.find(handle);
  if (it == ctx->textures.end()) return RTF_ERR_NOT_FOUND;
  if (n == 0) return RTF_OK;
  if (!uv || !rgba_out || n > 0xFFFFFFFFull) return RTF_ERR_INVALID;
  DevBuf d_in, d_out;
  int rc;
  if ((rc = upload(d_in, uv, 8 * n, ctx->stream)) || (rc = d_out.reserve(16 * n))) return rc;
  k_sample_texture<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dev_texture(it->second), (uint32_t)n, d_in.as<float>(),
                                                                         d_out.as<float4>());
  ctx->launches++;
  RTF_CUDA(cudaMemcpyAsync(rgba_out, d_out.p, 16 * n, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  d_in.release(), d_out.release();
  return RTF_OK;
}
int rtf_sample_environment(rtf_ctx* ctx, const rtf_env* env, size_t n, const float* pos_dir6, float* rgb_out) {
  CHECK_CTX(ctx);
  if (!env || n > 0xFFFFFFFFull) return RTF_ERR_INVALID;
  if (n == 0) return RTF_OK;
  if (!pos_dir6 || !rgb_out) return RTF_ERR_INVALID;
  DevScene sc = ctx->scene;
  if (env->environmental_map_id != 0) {  // usable without a committed scene: resolve the faces now
    const DevTexture* faces = nullptr;
    int rc0;
    if ((rc0 = resolve_environment_map(ctx, &faces))) return rc0;
    sc.env_faces = faces;
  }
  DevBuf d_in, d_out;
  int rc;
  if ((rc = upload(d_in, pos_dir6, 24 * n, ctx->stream)) || (rc = d_out.reserve(12 * n))) return rc;
  k_sample_environment<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(sc, *env, (uint32_t)n, d_in.as<float>(),
                                                                             d_out.as<float>());
  ctx->launches++;
  RTF_CUDA(cudaMemcpyAsync(rgb_out, d_out.p, 12 * n, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  d_in.release(), d_out.release();
  return RTF_OK;
}

int rtf_set_btf(rtf_ctx* ctx, uint64_t handle, const rtf_btf_desc* d) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  if (handle == 0 || !d) return RTF_ERR_INVALID;
  if (d->num_beta < 3 || (d->num_beta & 1) == 0 || d->num_alpha < 2 || d->num_theta < 2 || d->num_phi < 1 || d->pdf6_rows < 1 ||
      d->pdf6_cols < 1 || d->n_channels < 2 || d->pdf2d_slice_len < 2 || d->n_pdf1d < 1 || d->n_colors < 1 ||
      d->n_index_slices < 1 || d->n_pdf2d_color < 1 || d->n_pdf2d_lum < 1 || d->n_pdf2d < 1 || d->n_pdf3d < 1 || d->n_pdf4d < 1) {
    set_error("rtf_set_btf: bad table sizes (CompressedBTF.cpp:151-158 asserts)");
    return RTF_ERR_INVALID;
  }
  const void* ptrs[] = {d->pdf6_slices, d->pdf6_scales, d->pdf1d_basis, d->colors, d->index_ab, d->pdf2d_colors, d->pdf2d_slices,
                        d->pdf2d_scales, d->pdf2d_indices, d->pdf3d_scales, d->pdf3d_slices, d->pdf4d_scales, d->pdf4d_slices};
  for (const void* p : ptrs)
    if (!p) {
      set_error("rtf_set_btf: null table");
      return RTF_ERR_INVALID;
    }
  const size_t n6 = (size_t)d->pdf6_rows * d->pdf6_cols;
  bool ok = true;
  for (size_t i = 0; i < n6 && ok; i++) ok = d->pdf6_slices[i] >= 1 && d->pdf6_slices[i] <= d->n_pdf4d;  // 1-based, PDF6D.cuh:88
  ok = ok && indices_in_range(d->pdf4d_slices, (size_t)d->n_pdf4d * d->num_phi, 1, 0, d->n_pdf3d);
  ok = ok && indices_in_range(d->pdf3d_slices, (size_t)d->n_pdf3d * d->num_theta, 1, 0, d->n_pdf2d);
  ok = ok && indices_in_range(d->pdf2d_indices, (size_t)d->n_pdf2d * d->pdf2d_slice_len, d->pdf2d_slice_len, 0, d->n_pdf2d_lum);
  ok = ok && indices_in_range(d->pdf2d_indices + 1, (size_t)d->n_pdf2d * d->pdf2d_slice_len - 1, d->pdf2d_slice_len, 0,
                              d->n_pdf2d_color);
  ok = ok && indices_in_range(d->pdf2d_slices, (size_t)d->n_pdf2d_lum * d->num_alpha, 1, 0, d->n_pdf1d);
  ok = ok && indices_in_range(d->pdf2d_colors, (size_t)d->n_pdf2d_color * d->num_alpha, 1, 0, d->n_index_slices);
  ok = ok && indices_in_range(d->index_ab, (size_t)d->n_index_slices * d->num_beta, 1, 0, d->n_colors);
  if (!ok) {
    set_error("rtf_set_btf: an index table points outside the table it indexes");
    return RTF_ERR_INVALID;
  }
  cudaStreamSynchronize(ctx->stream);
  if (ctx->btfs.count(handle)) ctx->ready = false;  // the committed DevBTF table points at the arrays replaced below
  BTFTables& b = ctx->btfs[handle];
  b.desc = *d;
  // CompressedBTF.cpp:162-186: beta angles (degrees) of the 1-D slices
  std::vector<float> beta(d->num_beta);
  if (d->use_cos_beta) {
    for (int i = 0; i < d->num_beta; i++) {
      float s = -1.0f + 2.0f * i / (d->num_beta - 1);
      if (s > 1.0f) s = 1.0f;
      beta[i] = degreesf(asinp(s));
    }
    beta[0] = -90.f, beta[(d->num_beta - 1) / 2] = 0.f, beta[d->num_beta - 1] = 90.f;
  } else {
    const float step = 180.f / (d->num_beta - 1);
    for (int i = 0; i < d->num_beta; i++) beta[i] = i * step - 90.f;
    beta[(d->num_beta - 1) / 2] = 0.f, beta[d->num_beta - 1] = 90.0f;
  }
  cudaStream_t st = ctx->stream;
  int rc;
  if ((rc = upload(b.pdf6_slices, d->pdf6_slices, 4 * n6, st)) || (rc = upload(b.pdf6_scales, d->pdf6_scales, 4 * n6, st)) ||
      (rc = upload(b.pdf1d_basis, d->pdf1d_basis, 4 * (size_t)d->n_pdf1d * d->num_beta, st)) ||
      (rc = upload(b.colors, d->colors, 4 * (size_t)d->n_colors * d->n_channels, st)) ||
      (rc = upload(b.index_ab, d->index_ab, 4 * (size_t)d->n_index_slices * d->num_beta, st)) ||
      (rc = upload(b.pdf2d_colors, d->pdf2d_colors, 4 * (size_t)d->n_pdf2d_color * d->num_alpha, st)) ||
      (rc = upload(b.pdf2d_slices, d->pdf2d_slices, 4 * (size_t)d->n_pdf2d_lum * d->num_alpha, st)) ||
      (rc = upload(b.pdf2d_scales, d->pdf2d_scales, 4 * (size_t)d->n_pdf2d_lum * d->num_alpha, st)) ||
      (rc = upload(b.pdf2d_indices, d->pdf2d_indices, 4 * (size_t)d->n_pdf2d * d->pdf2d_slice_len, st)) ||
      (rc = upload(b.pdf3d_scales, d->pdf3d_scales, 4 * (size_t)d->n_pdf3d * d->num_theta, st)) ||
      (rc = upload(b.pdf3d_slices, d->pdf3d_slices, 4 * (size_t)d->n_pdf3d * d->num_theta, st)) ||
      (rc = upload(b.pdf4d_scales, d->pdf4d_scales, 4 * (size_t)d->n_pdf4d * d->num_phi, st)) ||
      (rc = upload(b.pdf4d_slices, d->pdf4d_slices, 4 * (size_t)d->n_pdf4d * d->num_phi, st)) ||
      (rc = upload(b.beta_angles, beta.data(), 4 * beta.size(), st)))
    return rc;
  RTF_CUDA(cudaStreamSynchronize(st));
  memset((void*)&b.desc.pdf6_slices, 0, sizeof(rtf_btf_desc) - offsetof(rtf_btf_desc, pdf6_slices));  // host pointers die here
  return RTF_OK;
}
int rtf_remove_btf(rtf_ctx* ctx, uint64_t handle) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  auto it = ctx->btfs.find(handle);
  if (it == ctx->btfs.end()) return RTF_ERR_NOT_FOUND;
  cudaStreamSynchronize(ctx->stream);
  it->second.release();
  ctx->btfs.erase(it);
  ctx->ready = false;
  return RTF_OK;
}
int rtf_set_material_btf(rtf_ctx* ctx, uint64_t material, uint64_t btf) {
  if (ctx) ctx->mat_dirty = true;
  CHECK_CTX(ctx);
  ctx->mat_links[material].btf = btf;
  return RTF_OK;
}
int rtf_btf_eval(rtf_ctx* ctx, uint64_t btf, size_t n, const float* tc, const float* view, const float* illum, const float* normal,
                 const float* tangent, float* rgb_out) {
  CHECK_CTX(ctx);
  auto it = ctx->btfs.find(btf);
  if (it == ctx->btfs.end()) return RTF_ERR_NOT_FOUND;
  if (n == 0) return RTF_OK;
  if (!tc || !view || !illum || !normal || !tangent || !rgb_out || n > 0xFFFFFFFFull) return RTF_ERR_INVALID;
  std::vector<float> in(14 * n);
  for (size_t i = 0; i < n; i++) {
    float* q = &in[14 * i];
    q[0] = tc[2 * i], q[1] = tc[2 * i + 1];
    for (int k = 0; k < 3; k++)
      q[2 + k] = view[3 * i + k], q[5 + k] = illum[3 * i + k], q[8 + k] = normal[3 * i + k], q[11 + k] = tangent[3 * i + k];
  }
  DevBuf d_in, d_out;
  int rc;
  if ((rc = upload(d_in, in.data(), 56 * n, ctx->stream)) || (rc = d_out.reserve(12 * n))) return rc;
  k_btf_eval<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(it->second.dev(), (uint32_t)n, d_in.as<float>(), d_out.as<float>());
  ctx->launches++;
  RTF_CUDA(cudaMemcpyAsync(rgb_out, d_out.p, 12 * n, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  d_in.release(), d_out.release();
  return RTF_OK;
}

// CompressedBTF::ImportFromFolder (CompressedBTF.cpp:74-390)
int rtf_btf_import_folder(rtf_ctx* ctx, uint64_t handle, const char* folder) {
  CHECK_CTX(ctx);
  if (handle == 0 || !folder) return RTF_ERR_INVALID;
  const std::string dir(folder);
  FILE* fp = fopen((dir + "/all_materialInfo.txt").c_str(), "r");
  if (!fp) {
    set_error("Failed to load BTF material: " + dir + "/all_materialInfo.txt");
    return RTF_ERR_NOT_FOUND;
  }
  rtf_btf_desc d;
  memset(&d, 0, sizeof(d));
  char line[1000], l1[1000], l2[1000], l3[1000], l4[1000];
  int load_materials, max_materials, flag_all, flag34, flag_compact, ncolour, nview, nillu, tile;
  int p[12], f1[10], f2[10], metric, ro = 0, co = 0, pr = 0, pc = 0;
  float eps[11], hdr_value = 1.0f;
  bool ok = fscanf(fp, "%999s %d %d %d %d %d", line, &load_materials, &max_materials, &flag_all, &flag34, &flag_compact) == 6;
  ok = ok && fscanf(fp, "%d %d %d %d", &ncolour, &nview, &nillu, &tile) == 4;
  for (int k = 0; k < 12 && ok; k++) ok = fscanf(fp, "%d", &p[k]) == 1;  // parameterisation, CompressedBTF.cpp:141-158
  for (int k = 0; k < 10 && ok; k++) ok = fscanf(fp, "%d", &f1[k]) == 1;  // search start flags (unused when rendering)
  for (int k = 0; k < 10 && ok; k++) ok = fscanf(fp, "%d", &f2[k]) == 1;
  ok = ok && fscanf(fp, "%d", &metric) == 1;
  for (int k = 0; k < 11 && ok; k++) ok = fscanf(fp, "%f", &eps[k]) == 1;
  if (!ok) {
    fclose(fp);
    set_error("all_materialInfo.txt is corrupted (basic parameters)");
    return RTF_ERR_INVALID;
  }
  if (!flag_all && load_materials != 1) {
    fclose(fp);
    set_error("Database for multiple materials are not supported!");  // CompressedBTF.cpp:245-248
    return RTF_ERR_INVALID;
  }
  std::string material;
  if (fscanf(fp, "%999s %999s %999s %999s %d %d %d %d %f", l1, l2, l3, l4, &ro, &co, &pr, &pc, &hdr_value) == 9) material = l1;
  fclose(fp);
  if (material.empty()) {
    set_error("all_materialInfo.txt names no material");
    return RTF_ERR_INVALID;
  }
  // <material>_materialInfo.txt, CompressedBTF.cpp:281-313
  fp = fopen((dir + "/" + material + "_materialInfo.txt").c_str(), "r");
  if (!fp) {
    set_error("Cannot open file " + dir + "/" + material + "_materialInfo.txt");
    return RTF_ERR_NOT_FOUND;
  }
  char name_m[1000];
  ok = fscanf(fp, "%999s %999s %999s %999s %d %d %d %d %f", name_m, l1, l2, l3, &ro, &co, &pr, &pc, &hdr_value) == 9;
  fclose(fp);
  if (!ok || material != name_m) {
    set_error("material info file is corrupted or names another material");
    return RTF_ERR_INVALID;
  }
  int hdr = 1;
  if (fabsf(hdr_value - 1.0f) < 1e-6f || fabsf(hdr_value) < 1e-6f) hdr = 0, hdr_value = 1.0f;
  d.use_cos_beta = p[0] ? 1 : 0, d.num_beta = p[1], d.num_alpha = p[3], d.num_theta = p[5], d.num_phi = p[7];
  d.code_btf_flag = p[11];
  d.hdr = hdr, d.hdr_value = hdr_value, d.multiplier = hdr ? hdr_value : 1.0f;  // CompressedBTF.cpp:380-384
  d.tex_coord_multiplier = 1.0f, d.gamma = 2.2f;                               // BTFBase.cuh:22-24
  d.rows_offset = ro, d.cols_offset = co;                                      // PDF6D::Init(pr, pc, ro, co), CompressedBTF.cpp:333
  std::vector<int32_t> pdf6_slices, index_ab, pdf2d_colors, pdf2d_slices, pdf2d_indices, pdf3d_slices, pdf4d_slices;
  std::vector<float> pdf6_scales, pdf1d, colors, pdf2d_scales, pdf3d_scales, pdf4d_scales;
  int r, c;
  const std::string mp = dir + "/" + material, ap = dir + "/all";
  if (!parse_table(mp + "_PDF6Dslices.txt", d.pdf6_rows, d.pdf6_cols, pdf6_slices)) return RTF_ERR_INVALID;
  if (!parse_table(mp + "_PDF6Dscale.txt", r, c, pdf6_scales) || r != d.pdf6_rows || c != d.pdf6_cols) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF1Dslice.txt", d.n_pdf1d, c, pdf1d) || c != d.num_beta) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_colors.txt", d.n_colors, d.n_channels, colors)) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_indexAB.txt", d.n_index_slices, c, index_ab) || c != d.num_beta) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF2Dcolours.txt", d.n_pdf2d_color, c, pdf2d_colors) || c != d.num_alpha) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF2Dslices.txt", d.n_pdf2d_lum, c, pdf2d_slices) || c != d.num_alpha) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF2Dscale.txt", r, c, pdf2d_scales) || r != d.n_pdf2d_lum || c != d.num_alpha) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF2Dindices.txt", d.n_pdf2d, d.pdf2d_slice_len, pdf2d_indices)) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF3Dscale.txt", d.n_pdf3d, c, pdf3d_scales) || c != d.num_theta) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF3Dslices.txt", r, c, pdf3d_slices) || r != d.n_pdf3d || c != d.num_theta) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF4Dscale.txt", d.n_pdf4d, c, pdf4d_scales) || c != d.num_phi) return RTF_ERR_INVALID;
  if (!parse_table(ap + "_PDF4Dslices.txt", r, c, pdf4d_slices) || r != d.n_pdf4d || c != d.num_phi) return RTF_ERR_INVALID;
  d.pdf6_slices = pdf6_slices.data(), d.pdf6_scales = pdf6_scales.data(), d.pdf1d_basis = pdf1d.data(), d.colors = colors.data();
  d.index_ab = index_ab.data(), d.pdf2d_colors = pdf2d_colors.data(), d.pdf2d_slices = pdf2d_slices.data();
  d.pdf2d_scales = pdf2d_scales.data(), d.pdf2d_indices = pdf2d_indices.data(), d.pdf3d_scales = pdf3d_scales.data();
  d.pdf3d_slices = pdf3d_slices.data(), d.pdf4d_scales = pdf4d_scales.data(), d.pdf4d_slices = pdf4d_slices.data();
  return rtf_set_btf(ctx, handle, &d);
}

}  // extern "C"
