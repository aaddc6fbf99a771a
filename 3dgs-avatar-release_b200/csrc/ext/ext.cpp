// ext.cpp -- the thin PyTorch C++ extension behind `diff_gaussian_rasterization._C` (north_star: "called from a thin PyTorch
// C++ extension").  Same three entry points with the same positional signatures as the reference's pybind11 module
// (submodules/diff-gaussian-rasterization/ext.cpp:15-19, rasterize_points.h:18-67, rasterize_points.cu:35-217):
// rasterize_gaussians, rasterize_gaussians_backward, mark_visible -- plus rasterize_gaussians_opacity (fused opacity pass).
// torch is used for device memory and the current stream only; all work happens in libgsplat_b200.so through its C ABI
// (include/gsplat_b200.h).  There is no CPU path: a non-CUDA tensor raises.
#include <ATen/cuda/CUDAContext.h>
#include <ATen/cuda/CUDAGraphsUtils.cuh>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <chrono>
#include <map>
#include <stdexcept>
#include <string>

#include "../../../include/gsplat_b200.h"

namespace {

struct DeviceState {            // per device: pair-capacity high-water mark, sort scratch, pinned pair-count word
  int64_t capacity = 0, graph_capacity = 0;
  torch::Tensor scratch, pinned;
  std::vector<torch::Tensor> graph_keepalive;  // a later regrow must not free what a captured graph points at
};
std::map<int, DeviceState> g_state;
int64_t g_forward_calls = 0, g_backward_calls = 0, g_regrows = 0;
constexpr double kGrow = 1.25;

DeviceState& dev_state(int dev) {
  auto& st = g_state[dev];
  if (!st.pinned.defined()) st.pinned = torch::empty({2}, torch::dtype(torch::kInt64).pinned_memory(true));
  return st;
}

[[noreturn]] void fail(const std::string& what, int status) {
  const char* msg = status == GS_ERR_BAD_ARG ? "bad argument" : status == GS_ERR_WORKSPACE_SMALL ? "workspace too small"
                  : status == GS_ERR_CUDA ? "CUDA error" : "error";
  std::string s = what + ": " + msg;
  if (status == GS_ERR_CUDA) s += " (cudaError " + std::to_string(gs_last_cuda_error()) + ")";
  throw std::runtime_error(s);
}

// device pointer of an optional FP32 tensor (the Python layer passes empty tensors for absent arguments, like the reference)
const float* fptr(const torch::Tensor& t, const char* name) {
  if (!t.defined() || t.numel() == 0) return nullptr;
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (there is no CPU path)");
  TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, " must be float32");
  TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
  return t.data_ptr<float>();
}
torch::Tensor contig(const torch::Tensor& t) { return (!t.defined() || t.is_contiguous()) ? t : t.contiguous(); }

using FwdResult = std::tuple<int64_t, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>;

// (num_rendered, color, opacity_image or undefined, radii, geomBuffer, binningBuffer, imgBuffer)
FwdResult forward_impl(const torch::Tensor& bg_, const torch::Tensor& means3D_, const torch::Tensor& colors_, const torch::Tensor& opacity_,
                       const torch::Tensor& scales_, const torch::Tensor& rotations_, double scale_modifier, const torch::Tensor& cov3D_,
                       const torch::Tensor& view_, const torch::Tensor& proj_, double tan_fovx, double tan_fovy, int64_t H, int64_t W,
                       const torch::Tensor& sh_, int64_t degree, const torch::Tensor& campos_, bool prefiltered, bool debug, bool want_opacity) {
  if (means3D_.dim() != 2 || means3D_.size(1) != 3) throw std::runtime_error("means3D must have dimensions (num_points, 3)");  // rasterize_points.cu:57-59
  TORCH_CHECK(means3D_.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
  const c10::cuda::CUDAGuard guard(means3D_.device());
  const int dev = means3D_.device().index();
  const int64_t P = means3D_.size(0);
  auto f32 = means3D_.options().dtype(torch::kFloat32);
  auto u8 = means3D_.options().dtype(torch::kUInt8);
  torch::Tensor op_img;
  if (want_opacity) op_img = torch::zeros({1, H, W}, f32);
  if (P == 0) {  // rasterize_points.cu:81
    auto z = torch::empty({0}, u8);
    return {0, torch::zeros({3, H, W}, f32), op_img, torch::zeros({0}, means3D_.options().dtype(torch::kInt32)), z, z.clone(), z.clone()};
  }
  auto means3D = contig(means3D_), colors = contig(colors_), opacity = contig(opacity_), scales = contig(scales_), rotations = contig(rotations_);
  auto cov3D = contig(cov3D_), sh = contig(sh_), bg = contig(bg_), view = contig(view_), proj = contig(proj_), campos = contig(campos_);
  auto out_color = torch::empty({3, H, W}, f32);
  auto radii = torch::empty({P}, means3D_.options().dtype(torch::kInt32));
  GsForwardArgs a{};
  a.P = (int32_t)P; a.D = (int32_t)degree; a.M = (sh.defined() && sh.numel() != 0) ? (int32_t)sh.size(1) : 0;
  a.width = (int32_t)W; a.height = (int32_t)H;
  a.tan_fovx = (float)tan_fovx; a.tan_fovy = (float)tan_fovy; a.scale_modifier = (float)scale_modifier;
  a.prefiltered = prefiltered; a.debug = debug;
  a.background = fptr(bg, "bg"); a.means3D = fptr(means3D, "means3D"); a.shs = fptr(sh, "sh"); a.colors_precomp = fptr(colors, "colors_precomp");
  a.opacities = fptr(opacity, "opacities"); a.scales = fptr(scales, "scales"); a.rotations = fptr(rotations, "rotations");
  a.cov3D_precomp = fptr(cov3D, "cov3D_precomp"); a.viewmatrix = fptr(view, "viewmatrix"); a.projmatrix = fptr(proj, "projmatrix");
  a.campos = fptr(campos, "campos");
  a.out_color = out_color.data_ptr<float>(); a.radii = radii.data_ptr<int32_t>();
  a.out_opacity = want_opacity ? op_img.data_ptr<float>() : nullptr;
  auto& st = dev_state(dev);
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(dev).stream();
  auto geom = torch::empty({(int64_t)gs_geom_bytes((int32_t)P)}, u8);
  auto img = torch::empty({(int64_t)gs_image_bytes((int32_t)W, (int32_t)H)}, u8);
  if (st.capacity == 0) st.capacity = std::max<int64_t>(4 * P, 1 << 18);
  const bool capturing = at::cuda::currentStreamCaptureStatus() != at::cuda::CaptureStatus::None;
  volatile int64_t* pn = st.pinned.data_ptr<int64_t>();
  while (true) {
    // capacities live on the 4096 lattice so that backward can recover them from the buffer size
    const int64_t cap = st.capacity = (st.capacity + 4095) / 4096 * 4096;
    auto lst = torch::empty({(int64_t)gs_list_bytes(cap, (int32_t)W, (int32_t)H)}, u8);
    const int64_t need = (int64_t)gs_scratch_bytes((int32_t)P, (int32_t)W, (int32_t)H, cap);
    if (!st.scratch.defined() || st.scratch.numel() < need || st.scratch.device() != means3D_.device())
      st.scratch = torch::empty({(int64_t)(need * 1.1) + 1024}, u8);
    pn[0] = -1; pn[1] = -1;
    const int status = gs_forward(&a, geom.data_ptr(), (size_t)geom.numel(), img.data_ptr(), (size_t)img.numel(), lst.data_ptr(),
                                  (size_t)lst.numel(), st.scratch.data_ptr(), (size_t)st.scratch.numel(), cap, (int64_t*)pn, stream);
    if (status != GS_OK) fail("gs_forward", status);
    g_forward_calls++;
    if (capturing) {
      // CUDA-graph capture: no host round trip.  The pair count lands in pinned memory on every replay; the caller checks it
      // with graph_overflow() after a replay (capacity is fixed inside a graph).
      st.graph_capacity = cap;
      st.graph_keepalive.push_back(st.scratch);
      return {-1, out_color, op_img, radii, geom, lst, img};
    }
    // The pair count N is produced on the device; the async copy into pinned memory lands while the blend kernel is still
    // running, so this wait overlaps GPU work (the reference blocks the pipeline instead, rasterizer_impl.cu:281).
    const auto t0 = std::chrono::steady_clock::now();
    uint64_t spins = 0;
    while (pn[0] == -1 || pn[1] == -1) {
      if ((++spins & 0xFFFF) == 0 && std::chrono::steady_clock::now() - t0 > std::chrono::seconds(30)) {
        cudaStreamSynchronize(stream);
        if (pn[0] == -1) throw std::runtime_error("gs_forward: the pair count never arrived (kernel failure?)");
      }
    }
    const int64_t N = pn[0], flags = pn[1];
    if ((flags >> 32) != 0) throw std::runtime_error("gs_forward: internal wait timeout (error flag " + std::to_string(flags >> 32) + ")");
    if (N <= cap) return {N, out_color, op_img, radii, geom, lst, img};
    st.capacity = (int64_t)(N * kGrow) + 4096;  // overflow: the image is invalid -> regrow and rerun
    g_regrows++;
  }
}

std::tuple<int64_t, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians(const torch::Tensor& bg, const torch::Tensor& means3D, const torch::Tensor& colors, const torch::Tensor& opacity,
                    const torch::Tensor& scales, const torch::Tensor& rotations, double scale_modifier, const torch::Tensor& cov3D_precomp,
                    const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, double tan_fovx, double tan_fovy, int64_t image_height,
                    int64_t image_width, const torch::Tensor& sh, int64_t degree, const torch::Tensor& campos, bool prefiltered, bool debug) {
  auto r = forward_impl(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                        tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, false);
  return {std::get<0>(r), std::get<1>(r), std::get<3>(r), std::get<4>(r), std::get<5>(r), std::get<6>(r)};
}

FwdResult rasterize_gaussians_opacity(const torch::Tensor& bg, const torch::Tensor& means3D, const torch::Tensor& colors,
                                      const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                                      double scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                                      const torch::Tensor& projmatrix, double tan_fovx, double tan_fovy, int64_t image_height,
                                      int64_t image_width, const torch::Tensor& sh, int64_t degree, const torch::Tensor& campos,
                                      bool prefiltered, bool debug) {
  return forward_impl(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                      tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, true);
}

// -> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations), rasterize_points.cu:196
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_backward(const torch::Tensor& bg_, const torch::Tensor& means3D_, const torch::Tensor& radii, const torch::Tensor& colors_,
                             const torch::Tensor& scales_, const torch::Tensor& rotations_, double scale_modifier, const torch::Tensor& cov3D_,
                             const torch::Tensor& view_, const torch::Tensor& proj_, double tan_fovx, double tan_fovy,
                             const torch::Tensor& dL_dout_color_, const torch::Tensor& sh_, int64_t degree, const torch::Tensor& campos_,
                             const torch::Tensor& geomBuffer, int64_t R, const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
                             bool debug, const c10::optional<torch::Tensor>& dL_dout_opacity_) {
  (void)R;
  const c10::cuda::CUDAGuard guard(means3D_.device());
  const int dev = means3D_.device().index();
  const int64_t P = means3D_.size(0), H = dL_dout_color_.size(1), W = dL_dout_color_.size(2);
  auto sh = contig(sh_);
  const int64_t M = (sh.defined() && sh.numel() != 0) ? sh.size(1) : 0;
  auto f32 = means3D_.options().dtype(torch::kFloat32);
  // one slab carved into the 8 gradient tensors of rasterize_points.cu:151-159.  NOT zero-filled: gs_backward writes every
  // element of every output (culled Gaussians get explicit zeros)
  const int64_t sizes[8] = {3, 3, 3, 1, 6, 3 * M, 3, 4};  // means3D, means2D, colors, opacity, cov3D, sh, scales, rot
  int64_t total = 0, off[8];
  for (int k = 0; k < 8; k++) { off[k] = total; total += (P * sizes[k] + 3) / 4 * 4; }
  auto flat = torch::empty({total}, f32);
  auto part = [&](int k) { return flat.narrow(0, off[k], P * sizes[k]); };
  auto dmeans3D = part(0).view({P, 3}), dmeans2D = part(1).view({P, 3}), dcolors = part(2).view({P, 3}), dopacity = part(3).view({P, 1});
  auto dcov3D = part(4).view({P, 6}), dsh = part(5).view({P, M, 3}), dscales = part(6).view({P, 3}), drot = part(7).view({P, 4});
  if (P == 0) return {dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot};
  auto means3D = contig(means3D_), colors = contig(colors_), scales = contig(scales_), rotations = contig(rotations_), cov3D = contig(cov3D_);
  auto bg = contig(bg_), view = contig(view_), proj = contig(proj_), campos = contig(campos_), dL = contig(dL_dout_color_);
  torch::Tensor dLo;
  if (dL_dout_opacity_.has_value() && dL_dout_opacity_->defined()) dLo = contig(*dL_dout_opacity_);
  GsBackwardArgs a{};
  a.P = (int32_t)P; a.D = (int32_t)degree; a.M = (int32_t)M; a.width = (int32_t)W; a.height = (int32_t)H;
  a.tan_fovx = (float)tan_fovx; a.tan_fovy = (float)tan_fovy; a.scale_modifier = (float)scale_modifier; a.debug = debug;
  a.background = fptr(bg, "bg"); a.means3D = fptr(means3D, "means3D"); a.shs = fptr(sh, "sh"); a.colors_precomp = fptr(colors, "colors_precomp");
  a.scales = fptr(scales, "scales"); a.rotations = fptr(rotations, "rotations"); a.cov3D_precomp = fptr(cov3D, "cov3D_precomp");
  a.viewmatrix = fptr(view, "viewmatrix"); a.projmatrix = fptr(proj, "projmatrix"); a.campos = fptr(campos, "campos");
  TORCH_CHECK(radii.is_cuda() && radii.scalar_type() == torch::kInt32 && radii.is_contiguous(), "radii must be a contiguous int32 CUDA tensor");
  a.radii = radii.data_ptr<int32_t>();
  a.dL_dpix = fptr(dL, "dL_dout_color");
  a.dL_dmean2D = dmeans2D.data_ptr<float>(); a.dL_dconic = nullptr; a.dL_dopacity = dopacity.data_ptr<float>();
  a.dL_dcolor = dcolors.data_ptr<float>(); a.dL_dmean3D = dmeans3D.data_ptr<float>(); a.dL_dcov3D = dcov3D.data_ptr<float>();
  a.dL_dsh = M ? dsh.data_ptr<float>() : nullptr; a.dL_dscale = dscales.data_ptr<float>(); a.dL_drot = drot.data_ptr<float>();
  a.dL_dopacity_pix = fptr(dLo, "dL_dout_opacity");
  const int64_t cap = gs_list_capacity((size_t)binningBuffer.numel(), (int32_t)W, (int32_t)H);
  const int status = gs_backward(&a, geomBuffer.data_ptr(), (size_t)geomBuffer.numel(), imageBuffer.data_ptr(), (size_t)imageBuffer.numel(),
                                 binningBuffer.data_ptr(), (size_t)binningBuffer.numel(), cap, at::cuda::getCurrentCUDAStream(dev).stream());
  if (status != GS_OK) fail("gs_backward", status);
  g_backward_calls++;
  return {dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot};
}

torch::Tensor mark_visible(const torch::Tensor& means3D_, const torch::Tensor& viewmatrix_, const torch::Tensor& projmatrix_) {
  TORCH_CHECK(means3D_.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
  const c10::cuda::CUDAGuard guard(means3D_.device());
  const int64_t P = means3D_.size(0);
  auto present = torch::zeros({P}, means3D_.options().dtype(torch::kBool));
  if (P) {
    auto means3D = contig(means3D_), view = contig(viewmatrix_), proj = contig(projmatrix_);
    const int status = gs_mark_visible((int32_t)P, fptr(means3D, "means3D"), fptr(view, "viewmatrix"), fptr(proj, "projmatrix"),
                                       (uint8_t*)present.data_ptr(), at::cuda::getCurrentCUDAStream(means3D_.device().index()).stream());
    if (status != GS_OK) fail("gs_mark_visible", status);
  }
  return present;
}

// ---- posing (avatar_pose.fused_pose) and the L1 loss (avatar_pose.loss.l1_loss): the same thin glue for the other two ops of the
// eager training step -- each ctypes call with its argument struct cost ~25 us of Python per call ----
using OptT = c10::optional<torch::Tensor>;

void fill_pose(PoseArgs& a, const torch::Tensor& xyz, const torch::Tensor& rot, const torch::Tensor& sc, const torch::Tensor& op,
               const torch::Tensor& dc, const torch::Tensor& rest, const torch::Tensor& w, const torch::Tensor& bones, const torch::Tensor& cam,
               int64_t sh_degree, double mod, bool cano, const torch::Tensor& vn) {
  if (w.dim() != 2 || w.size(1) != 24 || bones.numel() != 24 * 16) throw std::runtime_error("weights must be [P,24] and bone_transforms [24,4,4]");
  a.P = (int32_t)xyz.size(0); a.sh_degree = (int32_t)sh_degree; a.cano_view_dir = cano ? 1 : 0; a.scaling_modifier = (float)mod;
  a.M = 1 + ((rest.defined() && rest.numel()) ? (int32_t)rest.size(1) : 0);
  a.xyz = fptr(xyz, "xyz"); a.rotation = fptr(rot, "rotation"); a.scaling = fptr(sc, "scaling"); a.opacity = fptr(op, "opacity");
  a.features_dc = fptr(dc, "features_dc"); a.features_rest = fptr(rest, "features_rest"); a.weights = fptr(w, "weights");
  a.bone_transforms = fptr(bones, "bone_transforms"); a.camera_center = fptr(cam, "camera_center"); a.view_noise = fptr(vn, "view_noise");
}

// -> (out slab [P,13]: means 3 | cov 6 | colour 3 | opacity 1 as SoA slabs, T_fwd or undefined, rotation_bar or undefined)
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor>
pose_forward_t(const torch::Tensor& xyz, const torch::Tensor& rot, const torch::Tensor& sc, const torch::Tensor& op, const torch::Tensor& dc,
               const torch::Tensor& rest, const torch::Tensor& w, const torch::Tensor& bones, const torch::Tensor& cam, int64_t sh_degree,
               double mod, bool cano, bool want_aux, const OptT& view_noise) {
  TORCH_CHECK(xyz.is_cuda(), "fused_pose needs CUDA tensors (there is no CPU path)");
  const c10::cuda::CUDAGuard guard(xyz.device());
  const int64_t P = xyz.size(0);
  auto f32 = xyz.options().dtype(torch::kFloat32);
  auto out = torch::empty({P, 13}, f32);
  torch::Tensor T_fwd, Rbar, vn;
  if (want_aux) { T_fwd = torch::empty({P, 4, 4}, f32); Rbar = torch::empty({P, 3, 3}, f32); }
  if (view_noise.has_value() && view_noise->defined()) vn = *view_noise;
  if (P) {
    PoseArgs a{};
    fill_pose(a, xyz, rot, sc, op, dc, rest, w, bones, cam, sh_degree, mod, cano, vn);
    float* o = out.data_ptr<float>();
    a.means3D = o; a.cov3D = o + 3 * P; a.colors = o + 9 * P; a.opacity_out = o + 12 * P;
    a.T_fwd = want_aux ? T_fwd.data_ptr<float>() : nullptr;
    a.rotation_bar = want_aux ? Rbar.data_ptr<float>() : nullptr;
    const int status = pose_forward(&a, at::cuda::getCurrentCUDAStream(xyz.device().index()).stream());
    if (status != GS_OK) fail("pose_forward", status);
  }
  return {out, T_fwd, Rbar};
}

// -> (flat gradient slab, dL_dweights or undefined, dL_dbone_transforms or undefined); slab = xyz 3 | rot 4 | scale 3 | opacity 1 |
// dc 3 | rest 3 * nrest per Gaussian, attribute-major; `slab_buffer` (optional) is used when it is large enough
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor>
pose_backward_t(const torch::Tensor& xyz, const torch::Tensor& rot, const torch::Tensor& sc, const torch::Tensor& op, const torch::Tensor& dc,
                const torch::Tensor& rest, const torch::Tensor& w, const torch::Tensor& bones, const torch::Tensor& cam, int64_t sh_degree,
                double mod, bool cano, const OptT& view_noise, const OptT& g_means_, const OptT& g_cov_, const OptT& g_col_, const OptT& g_op_,
                bool need_w, bool need_b, const OptT& slab_buffer) {
  const c10::cuda::CUDAGuard guard(xyz.device());
  const int64_t P = xyz.size(0);
  auto f32 = xyz.options().dtype(torch::kFloat32);
  auto grad_or_zero = [&](const OptT& g, int64_t width) {
    if (g.has_value() && g->defined()) return contig(*g);
    return torch::zeros({P, width}, f32);
  };
  auto g_means = grad_or_zero(g_means_, 3), g_cov = grad_or_zero(g_cov_, 6), g_col = grad_or_zero(g_col_, 3), g_op = grad_or_zero(g_op_, 1);
  const int64_t nrest = (rest.defined() && rest.numel()) ? rest.size(1) : 0;
  const int64_t total = P * (3 + 4 + 3 + 1 + 3 + 3 * nrest);
  torch::Tensor flat;
  if (slab_buffer.has_value() && slab_buffer->defined() && slab_buffer->numel() >= total && slab_buffer->device() == xyz.device())
    flat = slab_buffer->narrow(0, 0, total);
  else
    flat = torch::empty({total}, f32);
  torch::Tensor gw, gb, vn;
  if (need_w) gw = torch::empty({P, 24}, f32);
  if (need_b) gb = torch::empty({24, 4, 4}, f32);
  if (view_noise.has_value() && view_noise->defined()) vn = *view_noise;
  if (P) {
    PoseBackwardArgs b{};
    fill_pose(b.fwd, xyz, rot, sc, op, dc, rest, w, bones, cam, sh_degree, mod, cano, vn);
    b.dL_dmeans3D = fptr(g_means, "dL_dmeans3D"); b.dL_dcov3D = fptr(g_cov, "dL_dcov3D"); b.dL_dcolors = fptr(g_col, "dL_dcolors");
    b.dL_dopacity = fptr(g_op, "dL_dopacity");
    float* o = flat.data_ptr<float>();
    b.dL_dxyz = o; b.dL_drotation = o + 3 * P; b.dL_dscaling = o + 7 * P; b.dL_dopacity_logit = o + 10 * P; b.dL_dfeatures_dc = o + 11 * P;
    b.dL_dfeatures_rest = nrest ? o + 14 * P : nullptr;
    b.dL_dweights = need_w ? gw.data_ptr<float>() : nullptr;
    b.dL_dbone_transforms = need_b ? gb.data_ptr<float>() : nullptr;
    const int status = pose_backward(&b, at::cuda::getCurrentCUDAStream(xyz.device().index()).stream());
    if (status != GS_OK) fail("pose_backward", status);
  }
  return {flat, gw, gb};
}

std::map<int, torch::Tensor> g_l1_ws;  // per device: partial sums + ticket of the fused L1 loss (zero on first use, self-resetting)

std::tuple<torch::Tensor, torch::Tensor> l1_forward_t(const torch::Tensor& image, const torch::Tensor& target) {
  TORCH_CHECK(image.is_cuda() && target.is_cuda(), "l1_loss needs CUDA tensors (there is no CPU path)");
  if (image.sizes() != target.sizes()) throw std::runtime_error("l1_loss: image and target shapes differ");
  const c10::cuda::CUDAGuard guard(image.device());
  const int dev = image.device().index();
  auto img = contig(image), gt = contig(target);
  auto& ws = g_l1_ws[dev];
  if (!ws.defined()) ws = torch::zeros({(int64_t)l1_loss_workspace_bytes()}, image.options().dtype(torch::kUInt8));
  auto loss = torch::empty({}, image.options().dtype(torch::kFloat32));
  auto unit = torch::empty_like(img);
  const int status = l1_loss_forward(fptr(img, "image"), fptr(gt, "target"), img.numel(), loss.data_ptr<float>(), unit.data_ptr<float>(),
                                     ws.data_ptr(), (size_t)ws.numel(), at::cuda::getCurrentCUDAStream(dev).stream());
  if (status != GS_OK) fail("l1_loss_forward", status);
  return {loss, unit};
}

torch::Tensor l1_backward_t(const torch::Tensor& unit, const torch::Tensor& g_) {
  const c10::cuda::CUDAGuard guard(unit.device());
  auto g = contig(g_);
  auto out = torch::empty_like(unit);
  const int status = l1_loss_backward(fptr(unit, "unit"), fptr(g, "dL_dloss"), unit.numel(), out.data_ptr<float>(),
                                      at::cuda::getCurrentCUDAStream(unit.device().index()).stream());
  if (status != GS_OK) fail("l1_loss_backward", status);
  return out;
}

// ---- capacity / graph bookkeeping (what the ctypes shim kept in _DeviceState) ----
int64_t get_capacity(int64_t dev) { return dev_state((int)dev).capacity; }
void set_capacity(int64_t dev, int64_t cap) { dev_state((int)dev).capacity = cap; }
std::tuple<int64_t, bool> graph_overflow(int64_t dev) {
  auto& st = dev_state((int)dev);
  const int64_t* pn = st.pinned.data_ptr<int64_t>();
  const int64_t N = pn[0], flags = pn[1];
  if ((flags >> 32) != 0) throw std::runtime_error("gs_forward: internal wait timeout (error flag " + std::to_string(flags >> 32) + ")");
  const bool over = N > (st.graph_capacity ? st.graph_capacity : st.capacity);
  if (over) { st.capacity = (int64_t)(N * kGrow) + 4096; g_regrows++; }
  return {N, over};
}
void release_cached() { g_state.clear(); g_l1_ws.clear(); }
std::tuple<int64_t, int64_t, int64_t> counters() { return {g_forward_calls, g_backward_calls, g_regrows}; }

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("rasterize_gaussians", &rasterize_gaussians);
  m.def("rasterize_gaussians_opacity", &rasterize_gaussians_opacity);
  m.def("rasterize_gaussians_backward", &rasterize_gaussians_backward, py::arg("bg"), py::arg("means3D"), py::arg("radii"), py::arg("colors"),
        py::arg("scales"), py::arg("rotations"), py::arg("scale_modifier"), py::arg("cov3D_precomp"), py::arg("viewmatrix"),
        py::arg("projmatrix"), py::arg("tan_fovx"), py::arg("tan_fovy"), py::arg("dL_dout_color"), py::arg("sh"), py::arg("degree"),
        py::arg("campos"), py::arg("geomBuffer"), py::arg("R"), py::arg("binningBuffer"), py::arg("imageBuffer"), py::arg("debug"),
        py::arg("dL_dout_opacity") = py::none());
  m.def("mark_visible", &mark_visible);
  m.def("pose_forward", &pose_forward_t);
  m.def("pose_backward", &pose_backward_t);
  m.def("l1_forward", &l1_forward_t);
  m.def("l1_backward", &l1_backward_t);
  m.def("get_capacity", &get_capacity);
  m.def("set_capacity", &set_capacity);
  m.def("graph_overflow", &graph_overflow);
  m.def("release_cached", &release_cached);
  m.def("counters", &counters);
}
