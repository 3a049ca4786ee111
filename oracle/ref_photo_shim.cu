// oracle/ref_photo_shim.cu — TEST INFRASTRUCTURE.  extern "C" access to the UNMODIFIED
// photometric registration of the reference's gsaligner submodule
// (submodules/gsaligner/gsaligner/registration/photometric_cuda.cu, photometric.cpp,
// tools/image_set*.c*), compiled where it lies by oracle/Makefile into
// oracle/_ref/libref_photo.so.  Only tests/ and bench.py's reference leg load it.
//
// The reference drives this code from GSAligner::setReference / setQuery / alignPhotometric
// (gsaligner/aligner.cu:181-250,318-342), which need torch; the few lines of glue they
// contain (which fields are set, in which order) are restated here without torch.  The
// reference translation unit is #included so that its kernels can be launched for the
// single-linearization probe below.
#include <registration/photometric_cuda.cu>

#include <cstring>
#include <memory>

namespace {

// aligner_utils.cu:3-16 (that file pulls in torch): depth > 1e-5 unmasks a pixel
__global__ void shim_init_image(const float *buf, photo::Image *img, int rows, int cols) {
  const int row = threadIdx.y + blockIdx.y * blockDim.y;
  const int col = threadIdx.x + blockIdx.x * blockDim.x;
  if (row >= rows || col >= cols) return;
  const float d = buf[row * cols + col];
  if (d > 1e-5) img->at<1>(row, col).setMasked(false);
  img->at<1>(row, col).setDepth(d);
}

struct RefPhoto {
  int H, W;
  photo::ImageAttributesPtr attributes;
  photo::SensorPtr sensor;
  photo::Image ref_image, query_image;
  photo::ImageSet ref_set, query_set;
  photo::PhotometricFac fact;
};

Eigen::Isometry3f pose_from_rowmajor(const float *T) {
  Eigen::Isometry3f X = Eigen::Isometry3f::Identity();
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) X.linear()(r, c) = T[4 * r + c];
    X.translation()(r) = T[4 * r + 3];
  }
  return X;
}

}  // namespace

extern "C" {

// GSAligner::GSAligner, aligner.h:94-121
void *ref_photo_create(int H, int W, float omega_depth, float depth_rejection_threshold,
                       float rho_kernel, float min_depth, float max_depth) {
  RefPhoto *p = new RefPhoto();
  p->H = H;
  p->W = W;
  p->attributes = std::make_shared<photo::ImageAttributes>();
  p->sensor = std::make_shared<photo::Sensor>();
  p->attributes->min_depth = min_depth;
  p->attributes->max_depth = max_depth;
  p->attributes->policies[photo::Intensity] = (photo::FilterPolicy)0;
  p->attributes->policies[photo::Depth] = (photo::FilterPolicy)2;
  p->attributes->thresholds[photo::Depth] = 0.5;
  p->attributes->policies[photo::Normal] = (photo::FilterPolicy)0;
  p->attributes->mask_grow_radius = 1.0;
  p->sensor->rows() = H;
  p->sensor->cols() = W;
  p->sensor->sensorOffset() = Eigen::Isometry3f::Identity();
  p->sensor->cameraType() = photo::CameraType::Spherical;
  p->fact.setOmegaDepth(omega_depth);
  p->fact.setDepthRejectionThreshold(depth_rejection_threshold);
  p->fact.setKernelChiThreshold(rho_kernel);
  return p;
}

void ref_photo_destroy(void *h) { delete static_cast<RefPhoto *>(h); }

// GSAligner::setReference, aligner.cu:181-218 (photometric half); K = row-major 3x3 on the host
int ref_photo_set_reference(void *h, const float *depth_dev, const float *K) {
  RefPhoto *p = static_cast<RefPhoto *>(h);
  p->ref_image = photo::Image(p->H, p->W);
  shim_init_image<<<dim3((p->W + 15) / 16, (p->H + 15) / 16), dim3(16, 16)>>>(
      depth_dev, p->ref_image.deviceInstance(), p->H, p->W);
  if (cudaDeviceSynchronize() != cudaSuccess) return 1;
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) p->sensor->cameraMatrix()(r, c) = K[3 * r + c];
  p->ref_set.matrix_ = p->ref_image;
  p->ref_set.attributes_ = p->attributes;
  p->ref_set.sensor_ = p->sensor;
  p->ref_set.growMaskDevice();
  p->ref_set.updateDerivativesDevice();
  p->fact.setFixedData(&p->ref_set);
  return cudaDeviceSynchronize() != cudaSuccess;
}

// GSAligner::setQuery, aligner.cu:219-258 (photometric half)
int ref_photo_set_query(void *h, const float *depth_dev) {
  RefPhoto *p = static_cast<RefPhoto *>(h);
  p->query_image = photo::Image(p->H, p->W);
  shim_init_image<<<dim3((p->W + 15) / 16, (p->H + 15) / 16), dim3(16, 16)>>>(
      depth_dev, p->query_image.deviceInstance(), p->H, p->W);
  if (cudaDeviceSynchronize() != cudaSuccess) return 1;
  p->query_set.matrix_ = p->query_image;
  p->query_set.attributes_ = p->attributes;
  p->query_set.sensor_ = p->sensor;
  p->fact.setMovingData(p->query_set);
  return cudaDeviceSynchronize() != cudaSuccess;
}

// GSAligner::alignPhotometric, aligner.cu:329-342.  T_in / T_out: row-major 4x4 on the host;
// stats = {residual, inlier ratio} as PhotometricFac::compute returns them.
int ref_photo_align(void *h, const float *T_in, int iterations, float *T_out, float *stats) {
  RefPhoto *p = static_cast<RefPhoto *>(h);
  p->fact.setTransform(pose_from_rowmajor(T_in));
  const auto st = p->fact.compute((size_t)iterations, false);
  const Eigen::Matrix4f M = p->fact.transform().matrix();
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) T_out[4 * r + c] = M(r, c);
  stats[0] = st.first;
  stats[1] = st.second;
  return cudaDeviceSynchronize() != cudaSuccess;
}

// One projection + linearization at pose T, WITHOUT the pose update: what
// PhotometricFac::linearize (photometric_cuda.cu:313-395) sums before it solves.  The entries
// the reference's linearize_kernel writes are copied back and added up here in double.
// out = upper triangle of H (21, row by row) | b (6) | chi | good | inliers  (30 doubles).
int ref_photo_linearize(void *h, const float *T, double *out) {
  RefPhoto *p = static_cast<RefPhoto *>(h);
  photo::PhotometricFac &f = p->fact;
  f.setTransform(pose_from_rowmajor(T));
  f.computeProjections();
  const size_t n = f.workspace_->size();
  photo::LinearSystemEntry *dev = nullptr;
  if (cudaMalloc((void **)&dev, sizeof(photo::LinearSystemEntry) * n) != cudaSuccess) return 1;
  photo::LinearSystemEntry zero;
  photo::fill_kernel<<<f.workspace_->nBlocks(), f.workspace_->nThreads()>>>(dev, zero, (uint32_t)n);
  photo::linearize_kernel<<<f.workspace_->nBlocks(), f.workspace_->nThreads()>>>(
      dev, f.workspace_->deviceInstance(), f.referenceData_->matrix_.deviceInstance(), f.SX_,
      f.neg2rotSX_, f.cameraMatrix_, f.camType_, f.kernelChiThreshold_, 1.0 / (double)n,
      std::sqrt(f.omegaIntensity_), std::sqrt(f.omegaDepth_), std::sqrt(f.omegaNormals_),
      f.depthRejectionThreshold_);
  if (cudaDeviceSynchronize() != cudaSuccess) return 1;
  photo::LinearSystemEntry *host = new photo::LinearSystemEntry[n];
  cudaMemcpy(host, dev, sizeof(photo::LinearSystemEntry) * n, cudaMemcpyDeviceToHost);
  for (int k = 0; k < 30; k++) out[k] = 0.0;
  for (size_t i = 0; i < n; i++) {
    for (int k = 0; k < 21; k++) out[k] += host[i].upperH(k);
    for (int k = 0; k < 6; k++) out[21 + k] += host[i].b(k);
    out[27] += (double)host[i].chi;
    out[28] += (double)host[i].is_good;
    out[29] += (double)host[i].is_inlier;
  }
  delete[] host;
  cudaFree(dev);
  return 0;
}

// The reference image after mask growing and derivative filtering (ImageSet::growMaskDevice,
// updateDerivativesDevice): depth | d/dcol | d/drow | masked, each H*W floats, to the host.
int ref_photo_reference_image(void *h, float *out) {
  RefPhoto *p = static_cast<RefPhoto *>(h);
  p->ref_set.matrix_.toHost();
  const size_t n = (size_t)p->H * p->W;
  for (size_t i = 0; i < n; i++) {
    const photo::ImageEntry &e = p->ref_set.matrix_.at(i);
    out[i] = e.value(1);
    out[n + i] = e.derivatives(1, 0);
    out[2 * n + i] = e.derivatives(1, 1);
    out[3 * n + i] = e.masked() ? 1.0f : 0.0f;
  }
  return 0;
}

}  // extern "C"
