// Host side above the C ABI: a headless mirror of the reference's plugin interface for the ray-casting path.
// Same class names, method names, argument meaning and bool-return error behaviour as the reference, so a renderer
// written against cppvolrend/volrenderbase.h:25-97 ports by changing its base-class include; GL objects are
// replaced by handles of libvrb200.so (include/vrb200.h), which is loaded with dlopen at first use — this library
// itself has no CUDA dependency and loads on a machine without a GPU (where every Init() then returns false).
#pragma once
#include <cstdint>
#include <fstream>
#include <string>
#include <vector>

#include "vrb200.h"

// ---- cppvolrend/utils/parameterspace.h:15-226 ------------------------------------------------------------
class ParameterRangeBase {
 public:
  explicit ParameterRangeBase(const std::string& name) : m_name(name) {}
  virtual ~ParameterRangeBase() {}
  virtual void Start() = 0;
  virtual void Incr() = 0;
  virtual bool End() const = 0;
  virtual int NumSteps() const = 0;
  virtual void SaveCurrentValue() = 0;
  virtual void RestoreCurrentValue() = 0;
  virtual std::string GetValueStr() const = 0;
  const std::string& GetName() const { return m_name; }
 protected:
  std::string m_name;
};

template <typename T>
class ParameterRangeNumeric : public ParameterRangeBase {
 public:
  ParameterRangeNumeric(const std::string& name, T* param, T start, T end, T incr);
  void Start() override { *m_curr = m_start; }
  void Incr() override { *m_curr += m_incr; }               // accumulates in T, like the reference (:122-126)
  bool End() const override { return *m_curr > m_end; }
  int NumSteps() const override;                            // 1 + ceil((end - start) / incr)   (:141-144)
  void SaveCurrentValue() override { m_saved = *m_curr; }
  void RestoreCurrentValue() override { *m_curr = m_saved; }
  std::string GetValueStr() const override { return std::to_string(*m_curr); }
 private:
  T m_start, m_end, m_incr, m_saved;
  T* m_curr;
};
using ParameterRangeFloat = ParameterRangeNumeric<float>;
using ParameterRangeDouble = ParameterRangeNumeric<double>;
using ParameterRangeInt = ParameterRangeNumeric<int>;

class ParameterSpace {
 public:
  ParameterSpace() {}
  virtual ~ParameterSpace() { ClearParameterDimensions(); }
  void AddParameterDimension(ParameterRangeBase* param);   // takes ownership
  void ClearParameterDimensions();
  int GetNumDimensions() const { return (int)m_dimensions.size(); }
  const std::string& GetDimensionName(int idx) const { return m_dimensions[idx]->GetName(); }
  const std::string GetDimensionValue(int idx) const { return m_dimensions[idx]->GetValueStr(); }
  int GetNumSamplePoints() const { return m_numsamples; }
  void StartEvaluation();
  void EndEvaluation();
  bool IncrEvaluation();                                    // odometer, first dimension fastest; false at the end
 private:
  std::vector<ParameterRangeBase*> m_dimensions;
  int m_numsamples = 0;
};

namespace vis {

enum class GRID_VOLUME_DATA_TYPE : unsigned int { STRUCTURED = 0, UNSTRUCTURED = 1, NONE_GRID = 2 };
enum class DataStorageSize : unsigned int { _8_BITS = 0, _16_BITS = 1, _NORMALIZED_F = 2, UNKNOWN = 3 };

struct vec3 { float x = 0, y = 0, z = 0; };

// ---- libs/vis_utils/camera.{h,cpp}: the subset Update() reads (eye, lookAt basis, fovy, aspect) -----------
class Camera {
 public:
  Camera();                                        // CameraData defaults: fovy 45, near 1, far 5000 (camera.cpp:17-54)
  void SetInitialConfig(vec3 eye, vec3 center, vec3 up);
  void UpdateAspectRatio(float w, float h) { m_aspect = w / h; }
  vec3 GetEye() const { return m_eye; }
  float GetFovY() const { return m_fovy; }
  void SetFovY(float deg) { m_fovy = deg; }
  float GetAspectRatio() const { return m_aspect; }
  void LookAt3(float out_m3[9]) const;             // mat3(glm::lookAt(eye, center, up)), column-major (camera.cpp:281-284)
  void FillUniforms(vrb_camera* out) const;        // what Update() binds: rc1prenderer.cpp:89-103
 private:
  vec3 m_eye, m_center, m_up;
  float m_fovy, m_aspect;
};

// ---- libs/volvis_utils/structuredgridvolume.{h,cpp} --------------------------------------------------------
class StructuredGridVolume {
 public:
  StructuredGridVolume(std::string name = "Unknown", unsigned w = 0, unsigned h = 0, unsigned d = 0);
  ~StructuredGridVolume();
  unsigned GetWidth() const { return m_w; }
  unsigned GetHeight() const { return m_h; }
  unsigned GetDepth() const { return m_d; }
  double GetScaleX() const { return m_sx; }
  double GetScaleY() const { return m_sy; }
  double GetScaleZ() const { return m_sz; }
  void SetScale(double sx, double sy, double sz) { m_sx = sx; m_sy = sy; m_sz = sz; }
  void SetArrayData(void* data, DataStorageSize dss);          // takes ownership (malloc'ed)
  void* GetArrayData() const { return m_data; }
  double GetNormalizedSample(int x, int y, int z) const;       // structuredgridvolume.cpp:121-151
  const std::string& GetName() const { return m_name; }
  DataStorageSize m_data_storage_size = DataStorageSize::UNKNOWN;
 private:
  std::string m_name;
  unsigned m_w, m_h, m_d;
  double m_sx = 1, m_sy = 1, m_sz = 1;
  void* m_data = nullptr;
};

// ---- libs/volvis_utils/transferfunction1d.{h,cpp} ----------------------------------------------------------
class TransferFunction1D {
 public:
  explicit TransferFunction1D(int max_value = 255) : max_density(max_value) {}
  void SetExtinctionCoefficientInput(bool s) { extinction_coef_type = s; }
  void AddRGBControlPoint(double r, double g, double b, int isovalue);
  void AddAlphaControlPoint(double alpha, int isovalue);
  void Build();                                                 // BuildLinear, transferfunction1d.cpp:319-358
  const std::vector<double>& Table() const { return m_table; }  // (max_density+1) x {r,g,b,a}
  int max_density;
  bool extinction_coef_type = false;
 private:
  struct Cp { float v[4]; int iso; };
  std::vector<Cp> m_rgb, m_alpha;
  std::vector<double> m_table;
};

// ---- libs/volvis_utils/reader.cpp: .raw (:162-225 + :62-98), .syn (:283-371), .pvm (:100-160 + file_utils/pvm.cpp), .nrrd / .nhrd
// (:373-496), .dat (:498-628), .tf1d (:744-814) ---------------------------------------------------------------
class VolumeReader {
 public:
  StructuredGridVolume* ReadVolume(const std::string& path);   // dispatch on extension; nullptr on failure
  StructuredGridVolume* readraw(const std::string& path);
  StructuredGridVolume* readsyn(const std::string& path);
  StructuredGridVolume* readpvm(const std::string& path);      // PVM / PVM2 / PVM3, plain or DDS v3d / v3e compressed, 8 or 16 bits
  StructuredGridVolume* readpvmold(const std::string& path);   // the same container through PvmOld: min-max stretched, normalised floats
  StructuredGridVolume* readnrrd(const std::string& path);     // detached header: type, sizes, data file
  StructuredGridVolume* readnhrd(const std::string& path);
  StructuredGridVolume* readdat(const std::string& path);      // ObjectFileName, Resolution, SliceThickness, Format
};
class TransferFunctionReader {
 public:
  TransferFunction1D* ReadTransferFunction(const std::string& path);   // .tf1d; nullptr on failure
};

// ---- libs/volvis_utils/renderingparameters.cpp:17-32: Blinn-Phong + first light ----------------------------
class RenderingParameters {
 public:
  RenderingParameters();
  void FillUniforms(vrb_light* out) const;
  float ka, kd, ks, shininess;
  vec3 light_position, light_specular;
};

// ---- libs/volvis_utils/datamanager.h:82-99: what renderers read through m_ext_data_manager -----------------
// Owns the vrb context: SetCurrentVolume / SetCurrentTransferFunction replace GenerateRTexture and
// GenerateTexture_1D_* (the device textures live in libvrb200.so).
// libs/volvis_utils/datamanager.h: how the gradient texture of the Blinn-Phong path is computed
enum class STRUCTURED_GRADIENT_TYPE { NONE_GRADIENT = 0, SOBEL_FELDMAN_FILTER = 1, FINITE_DIFERENCES = 2, COMPUTE_SHADER_SOBEL = 3 };   // datamanager.h:64-69 (values = vrb_volume_gradient models)

class DataManager {
 public:
  DataManager();
  ~DataManager();
  bool InitDevice(int device);                                           // false without a B200 (message in LastError)
  bool SetCurrentVolume(StructuredGridVolume* vol);                      // takes ownership
  bool SetCurrentTransferFunction(TransferFunction1D* tf);               // takes ownership
  StructuredGridVolume* GetCurrentStructuredVolume() { return m_vol; }
  TransferFunction1D* GetCurrentTransferFunction() { return m_tf; }
  bool HasVolumeTexture() const { return m_vol_uploaded; }               // GetCurrentVolumeTexture() != nullptr
  bool HasGradientTexture() const { return m_has_gradient; }             // GetCurrentGradientTexture() != nullptr
  bool SetCurrentGradientGenerator(STRUCTURED_GRADIENT_TYPE t);          // + GenerateStructuredGradientTexture (datamanager.cpp:332-352)
  bool GenerateStructuredGradientTexture();
  STRUCTURED_GRADIENT_TYPE curr_gradient_comp_model = STRUCTURED_GRADIENT_TYPE::NONE_GRADIENT;   // datamanager.cpp:27
  GRID_VOLUME_DATA_TYPE GetInputVolumeDataType() const { return GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  vrb_ctx* Context() { return m_ctx; }
  const std::string& LastError() const { return m_err; }
 private:
  vrb_ctx* m_ctx = nullptr;
  StructuredGridVolume* m_vol = nullptr;
  TransferFunction1D* m_tf = nullptr;
  bool m_vol_uploaded = false, m_has_gradient = false;
  std::string m_err;
};

}  // namespace vis
