// veles_b200 native runtime — inference units backed by the sm_100a kernel library (same kernels the Python
// framework trains with).  Registered by class UUID + name (reference fixture: libVeles/tests/workflow_files/
// contents.json — All2AllTanh b3a2bd5c-..., All2AllSoftmax 420219fc-...).
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "../../include/veles_b200.h"
#include "veles/unit_factory.h"

namespace veles {

static void vb_check(int rc, const char* what) {
  if (rc) throw std::runtime_error(std::string(what) + ": " + vb_last_error());
}
#define VU_CUDA(expr)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess) throw std::runtime_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

static uint16_t f2bf(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7F800000u) == 0x7F800000u && (u & 0x7FFFFFu)) return static_cast<uint16_t>((u >> 16) | 0x40);
  u += 0x7FFFu + ((u >> 16) & 1);
  return static_cast<uint16_t>(u >> 16);
}

static int activation_code(const std::string& mode) {
  if (mode == "ACTIVATION_TANH") return 1;
  if (mode == "ACTIVATION_SIGMOID") return 2;
  if (mode == "ACTIVATION_RELU") return 3;
  if (mode == "ACTIVATION_STRICT_RELU") return 4;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
class WeightedUnit : public Unit {
 public:
  ~WeightedUnit() override {
    if (w_dev_) cudaFree(w_dev_);
    if (b_dev_) cudaFree(b_dev_);
  }
  void SetParameter(const std::string& name, const Property& v) override {
    if (name == "weights") weights_ = v.array;
    else if (name == "bias") bias_ = v.array;
    else if (name == "include_bias") include_bias_ = v.b;
    else if (name == "weights_transposed") transposed_ = v.b;
    else if (name == "activation_mode") act_ = activation_code(v.str);
    else SetExtra(name, v);
  }

 protected:
  virtual void SetExtra(const std::string&, const Property&) {}
  // uploads weights as bf16 [n_out, pitch] (row pitch multiple of 8) and bias as fp32
  void Upload(int64_t n_out, int64_t n_in) {
    if (!weights_) throw std::runtime_error(std::string(Name()) + ": weights were not set");
    if (weights_->size() != n_out * n_in) throw std::runtime_error(std::string(Name()) + ": weights size mismatch");
    pitch_ = (n_in + 7) / 8 * 8;
    std::vector<uint16_t> host(static_cast<size_t>(n_out) * pitch_, 0);
    for (int64_t o = 0; o < n_out; ++o)
      for (int64_t i = 0; i < n_in; ++i) {
        const float w = transposed_ ? weights_->data[i * n_out + o] : weights_->data[o * n_in + i];
        host[o * pitch_ + i] = f2bf(w);
      }
    VU_CUDA(cudaMalloc(&w_dev_, host.size() * 2));
    VU_CUDA(cudaMemcpy(w_dev_, host.data(), host.size() * 2, cudaMemcpyHostToDevice));
    if (include_bias_ && bias_) {
      if (bias_->size() != n_out) throw std::runtime_error(std::string(Name()) + ": bias size mismatch");
      VU_CUDA(cudaMalloc(&b_dev_, n_out * 4));
      VU_CUDA(cudaMemcpy(b_dev_, bias_->data.data(), n_out * 4, cudaMemcpyHostToDevice));
    }
  }
  std::shared_ptr<NumpyArray> weights_, bias_;
  bool include_bias_ = true, transposed_ = false;
  int act_ = 0;
  void* w_dev_ = nullptr;
  void* b_dev_ = nullptr;
  int64_t pitch_ = 0;
};

// ---------------------------------------------------------------------------------------------------------
template <int ACT, bool SOFTMAX>
class All2AllUnit : public WeightedUnit {
 public:
  Shape OutputShape(const Shape&) const override {
    if (!weights_) throw std::runtime_error("All2All: weights were not set");
    return Shape{transposed_ ? weights_->shape[1] : weights_->shape[0]};
  }
  size_t ScratchBytes(int batch, const Shape& in) const override {
    // staging for inputs whose sample size is not a multiple of 8 (TMA pitch), plus fp32 logits for softmax
    const int64_t n_in = shape_size(in), n_out = OutputShape(in)[0];
    size_t s = (n_in % 8) ? static_cast<size_t>(batch) * ((n_in + 7) / 8 * 8) * 2 : 0;
    if (SOFTMAX) s += static_cast<size_t>(batch) * ((n_out + 3) / 4 * 4) * 4 + 256;
    return s;
  }
  void Initialize(int, const Shape& in) override { Upload(OutputShape(in)[0], shape_size(in)); }
  void Execute(const void* in, void* out, void* scratch, int batch, const Shape& ishape, cudaStream_t s) override {
    const int64_t n_in = shape_size(ishape), n_out = OutputShape(ishape)[0];
    const void* a = in;
    long long lda = n_in;
    char* sc = static_cast<char*>(scratch);
    if (n_in % 8) {
      lda = (n_in + 7) / 8 * 8;
      vb_check(vb_copy_2d(in, 1, n_in, sc, 1, lda, batch, static_cast<int>(n_in), s), "vb_copy_2d");
      a = sc;
      sc += (static_cast<size_t>(batch) * lda * 2 + 255) / 256 * 256;
    }
    const int act = act_ ? act_ : ACT;
    if (!SOFTMAX) {
      vb_check(vb_gemm_bf16(a, lda, 0, w_dev_, pitch_, 0, out, n_out, batch, static_cast<int>(n_out),
                            static_cast<int>(n_in), static_cast<const float*>(b_dev_), act, 0, 0, nullptr, 0, 0, 1.0f,
                            1, 0, s), "vb_gemm_bf16");
    } else {
      float* logits = reinterpret_cast<float*>(sc);
      const long long ldl = (n_out + 3) / 4 * 4;
      vb_check(vb_gemm_bf16(a, lda, 0, w_dev_, pitch_, 0, logits, ldl, batch, static_cast<int>(n_out),
                            static_cast<int>(n_in), static_cast<const float*>(b_dev_), 0, 1, 0, nullptr, 0, 0, 1.0f,
                            1, 0, s), "vb_gemm_bf16");
      vb_check(vb_softmax_eval(logits, 0, ldl, out, n_out, nullptr, 0, 1, nullptr, nullptr, batch,
                               static_cast<int>(n_out), 1.0f, nullptr, nullptr, nullptr, nullptr, s), "vb_softmax_eval");
    }
  }
};

#define VELES_A2A(cls, ACT, SM, uuid, name)                         \
  class cls : public All2AllUnit<ACT, SM> {                          \
   public:                                                           \
    const char* Uuid() const override { return uuid; }               \
    const char* Name() const override { return name; }               \
  };                                                                 \
  VELES_REGISTER_UNIT(cls, uuid, name)

VELES_A2A(All2All, 0, false, "58a5eadf-ae1e-498f-bf35-7d93939c4c86", "All2All");
VELES_A2A(All2AllTanh, 1, false, "b3a2bd5c-3c01-46ef-978a-fef22e008f31", "All2AllTanh");
VELES_A2A(All2AllSigmoid, 2, false, "a27974ec-1764-4944-925d-4862de237881", "All2AllSigmoid");
VELES_A2A(All2AllRELU, 3, false, "5b7f36d8-f8c8-4eb7-8af3-75eb3cfca3fe", "All2AllRELU");
VELES_A2A(All2AllStrictRELU, 4, false, "fe63baf0-4fe4-4cf3-bafb-ef1215bf27a8", "All2AllStrictRELU");
VELES_A2A(All2AllSoftmax, 0, true, "420219fc-3e1a-45b1-87f8-aaa0c1540de4", "All2AllSoftmax");

// ---------------------------------------------------------------------------------------------------------
template <int ACT>
class ConvUnit : public WeightedUnit {
 public:
  Shape OutputShape(const Shape& in) const override {
    const Shape s = In3(in);
    return Shape{(s[0] + pad_[1] + pad_[3] - ky_) / slide_[1] + 1, (s[1] + pad_[0] + pad_[2] - kx_) / slide_[0] + 1,
                 n_kernels_};
  }
  // Layers whose per-group channel count allows 16-byte channel vectors run as an implicit GEMM (TMA im2col-mode
  // operand loads inside vb_conv_gemm_bf16): no column matrix, hence no scratch.
  bool Implicit(const Shape& in) const {
    const Shape s = In3(in);
    return (s[2] / groups_) % 8 == 0 && (n_kernels_ / groups_) % 8 == 0;
  }
  size_t ScratchBytes(int batch, const Shape& in) const override {
    if (Implicit(in)) return 0;
    const Shape o = OutputShape(in);
    return static_cast<size_t>(batch) * o[0] * o[1] * ColsPitch(in) * 2;
  }
  void Initialize(int, const Shape& in) override {
    const Shape s = In3(in);
    Upload(n_kernels_, static_cast<int64_t>(ky_) * kx_ * s[2] / groups_);
  }
  void Execute(const void* in, void* out, void* scratch, int batch, const Shape& ishape, cudaStream_t st) override {
    const Shape s = In3(ishape), o = OutputShape(ishape);
    const long long ld = ColsPitch(ishape);
    if (Implicit(ishape)) {
      const int cg = static_cast<int>(s[2] / groups_), nkg_i = static_cast<int>(n_kernels_ / groups_);
      const int M_i = static_cast<int>(batch * o[0] * o[1]);
      const int act_i = act_ ? act_ : ACT;
      for (int g = 0; g < groups_; ++g) {
        const char* x = static_cast<const char*>(in) + static_cast<size_t>(g) * cg * 2;
        const char* w = static_cast<const char*>(w_dev_) + static_cast<size_t>(g) * nkg_i * pitch_ * 2;
        char* d = static_cast<char*>(out) + static_cast<size_t>(g) * nkg_i * 2;
        const float* b = b_dev_ ? static_cast<const float*>(b_dev_) + g * nkg_i : nullptr;
        vb_check(vb_conv_gemm_bf16(1, x, batch, static_cast<int>(s[0]), static_cast<int>(s[1]), s[2], cg, ky_, kx_,
                                   slide_[1], slide_[0], -pad_[1], -pad_[0], static_cast<int>(o[0]),
                                   static_cast<int>(o[1]), w, pitch_, nkg_i, cg, d, n_kernels_, M_i, nkg_i, b, act_i, 0,
                                   0, nullptr, 0, 0, 1.0f, 1, 0, st),
                 "vb_conv_gemm_bf16");
      }
      return;
    }
    vb_check(vb_im2col(in, scratch, batch, static_cast<int>(s[0]), static_cast<int>(s[1]), static_cast<int>(s[2]),
                       static_cast<int>(o[0]), static_cast<int>(o[1]), ky_, kx_, slide_[1], slide_[0], pad_[1], pad_[0],
                       groups_, ld, st), "vb_im2col");
    const int M = static_cast<int>(batch * o[0] * o[1]);
    const int kg = static_cast<int>(static_cast<int64_t>(ky_) * kx_ * s[2] / groups_);
    const int nkg = static_cast<int>(n_kernels_ / groups_);
    const int act = act_ ? act_ : ACT;
    for (int g = 0; g < groups_; ++g) {
      const char* a = static_cast<const char*>(scratch) + static_cast<size_t>(g) * kg * 2;
      const char* w = static_cast<const char*>(w_dev_) + static_cast<size_t>(g) * nkg * pitch_ * 2;
      char* d = static_cast<char*>(out) + static_cast<size_t>(g) * nkg * 2;
      const float* b = b_dev_ ? static_cast<const float*>(b_dev_) + g * nkg : nullptr;
      vb_check(vb_gemm_bf16(a, ld, 0, w, pitch_, 0, d, n_kernels_, M, nkg, kg, b, act, 0, 0, nullptr, 0, 0, 1.0f, 1, 0,
                            st), "vb_gemm_bf16");
    }
  }

 protected:
  void SetExtra(const std::string& name, const Property& v) override {
    if (name == "n_kernels") n_kernels_ = static_cast<int64_t>(v.num);
    else if (name == "kx") kx_ = static_cast<int>(v.num);
    else if (name == "ky") ky_ = static_cast<int>(v.num);
    else if (name == "grouping") groups_ = static_cast<int>(v.num);
    else if (name == "padding" && v.list.size() == 4) for (int i = 0; i < 4; ++i) pad_[i] = static_cast<int>(v.list[i]);
    else if (name == "sliding" && v.list.size() == 2) for (int i = 0; i < 2; ++i) slide_[i] = static_cast<int>(v.list[i]);
  }
  static Shape In3(const Shape& in) {
    if (in.size() == 3) return in;
    if (in.size() == 2) return Shape{in[0], in[1], 1};
    throw std::runtime_error("Conv: input must be HWC");
  }
  long long ColsPitch(const Shape& in) const {
    const Shape s = In3(in);
    return (static_cast<long long>(ky_) * kx_ * s[2] + 7) / 8 * 8;
  }
  int64_t n_kernels_ = 0;
  int kx_ = 1, ky_ = 1, groups_ = 1;
  int pad_[4] = {0, 0, 0, 0};   // left, top, right, bottom
  int slide_[2] = {1, 1};       // x, y
};

#define VELES_CONV(cls, ACT, uuid, name)                \
  class cls : public ConvUnit<ACT> {                     \
   public:                                               \
    const char* Uuid() const override { return uuid; }   \
    const char* Name() const override { return name; }   \
  };                                                     \
  VELES_REGISTER_UNIT(cls, uuid, name)

VELES_CONV(Conv, 0, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0001", "Conv");
VELES_CONV(ConvTanh, 1, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0002", "ConvTanh");
VELES_CONV(ConvSigmoid, 2, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0003", "ConvSigmoid");
VELES_CONV(ConvRELU, 3, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0004", "ConvRELU");
VELES_CONV(ConvStrictRELU, 4, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0005", "ConvStrictRELU");

// ---------------------------------------------------------------------------------------------------------
class PoolingUnit : public Unit {
 public:
  explicit PoolingUnit(int kind) : kind_(kind) {}
  void SetParameter(const std::string& name, const Property& v) override {
    if (name == "kx") kx_ = static_cast<int>(v.num);
    else if (name == "ky") ky_ = static_cast<int>(v.num);
    else if (name == "sliding" && v.list.size() == 2) { sx_ = static_cast<int>(v.list[0]); sy_ = static_cast<int>(v.list[1]); }
  }
  static int64_t OutSize(int64_t size, int k, int s) { return size <= k ? 1 : (size - k + s - 1) / s + 1; }
  Shape OutputShape(const Shape& in) const override {
    return Shape{OutSize(in[0], ky_, sy_), OutSize(in[1], kx_, sx_), in.size() > 2 ? in[2] : 1};
  }
  void Execute(const void* in, void* out, void*, int batch, const Shape& s, cudaStream_t st) override {
    const Shape o = OutputShape(s);
    vb_check(vb_pool_forward(kind_, in, out, nullptr, 1, batch, static_cast<int>(s[0]), static_cast<int>(s[1]),
                             static_cast<int>(o[2]), static_cast<int>(o[0]), static_cast<int>(o[1]), ky_, kx_, sy_, sx_,
                             nullptr, st), "vb_pool_forward");
  }

 private:
  int kind_, kx_ = 2, ky_ = 2, sx_ = 2, sy_ = 2;
};
class MaxPooling : public PoolingUnit {
 public:
  MaxPooling() : PoolingUnit(0) {}
  const char* Uuid() const override { return "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0011"; }
  const char* Name() const override { return "MaxPooling"; }
};
class MaxAbsPooling : public PoolingUnit {
 public:
  MaxAbsPooling() : PoolingUnit(1) {}
  const char* Uuid() const override { return "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0012"; }
  const char* Name() const override { return "MaxAbsPooling"; }
};
class AvgPooling : public PoolingUnit {
 public:
  AvgPooling() : PoolingUnit(2) {}
  const char* Uuid() const override { return "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0013"; }
  const char* Name() const override { return "AvgPooling"; }
};
VELES_REGISTER_UNIT(MaxPooling, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0011", "MaxPooling");
VELES_REGISTER_UNIT(MaxAbsPooling, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0012", "MaxAbsPooling");
VELES_REGISTER_UNIT(AvgPooling, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0013", "AvgPooling");

// ---------------------------------------------------------------------------------------------------------
// Recurrent layers (veles_b200/znicz/recurrent.py): the concatenated operand [x_t | h_{t-1}] of every step lives
// time-major in one bf16 scratch buffer; a step is one GEMM (bias in the epilogue) + one cell kernel that writes h_t
// into the next step's slot and into the output rows.  GATES = 1: Elman RNN (tanh), 4: LSTM (i f g o).
template <int GATES>
class RecurrentUnit : public WeightedUnit {
 public:
  Shape OutputShape(const Shape& in) const override {
    int64_t T, D;
    Geometry(in, &T, &D);
    if (return_sequences_) return Shape{T, hidden_};
    return Shape{hidden_};
  }
  size_t ScratchBytes(int batch, const Shape& in) const override {
    int64_t T, D;
    Geometry(in, &T, &D);
    const Layout l = MakeLayout(batch, T, D);
    return l.total;
  }
  void Initialize(int, const Shape& in) override {
    int64_t T, D;
    Geometry(in, &T, &D);
    Upload(GATES * hidden_, D + hidden_);
  }
  void Execute(const void* in, void* out, void* scratch, int batch, const Shape& ishape, cudaStream_t s) override {
    int64_t T, D;
    Geometry(ishape, &T, &D);
    const int64_t H = hidden_;
    const Layout l = MakeLayout(batch, T, D);
    char* base = static_cast<char*>(scratch);
    uint16_t* xh = reinterpret_cast<uint16_t*>(base + l.xh);
    float* z = reinterpret_cast<float*>(base + l.z);
    float* state = reinterpret_cast<float*>(base + l.state);      // LSTM: gates | c (ping) | c (pong); RNN: h
    // h_{-1} = 0: the h slot of step 0 is the only one no cell kernel writes
    VU_CUDA(cudaMemset2DAsync(xh + D, l.pitch * 2, 0, H * 2, batch, s));
    const uint16_t* x = static_cast<const uint16_t*>(in);
    for (int64_t t = 0; t < T; ++t) {
      uint16_t* xh_t = xh + t * batch * l.pitch;
      vb_check(vb_copy_2d(x + t * D, 1, T * D, xh_t, 1, l.pitch, batch, static_cast<int>(D), s), "vb_copy_2d");
    }
    uint16_t* o = static_cast<uint16_t*>(out);
    const long long ld_out = return_sequences_ ? T * H : H;
    for (int64_t t = 0; t < T; ++t) {
      uint16_t* xh_t = xh + t * batch * l.pitch;
      vb_check(vb_gemm_bf16(xh_t, l.pitch, 0, w_dev_, pitch_, 0, z, l.zpitch, batch, static_cast<int>(GATES * H),
                            static_cast<int>(D + H), static_cast<const float*>(b_dev_), 0, 1, 0, nullptr, 0, 0, 1.0f,
                            1, 0, s), "vb_gemm_bf16");
      void* h_next = (t + 1 < T) ? static_cast<void*>(xh + (t + 1) * batch * l.pitch + D) : nullptr;
      void* orow = return_sequences_ ? static_cast<void*>(o + t * H) : (t + 1 == T ? static_cast<void*>(o) : nullptr);
      if (GATES == 4) {
        float* gates = state;
        float* c0 = state + static_cast<size_t>(batch) * 4 * H;
        float* c1 = c0 + static_cast<size_t>(batch) * H;
        float* c_prev = (t & 1) ? c0 : c1;
        float* c_cur = (t & 1) ? c1 : c0;
        vb_check(vb_lstm_cell_forward(z, l.zpitch, t ? c_prev : nullptr, c_cur, gates, 4 * H, h_next, l.pitch, orow, 1,
                                      ld_out, batch, static_cast<int>(H), s), "vb_lstm_cell_forward");
      } else {
        vb_check(vb_rnn_cell_forward(z, l.zpitch, state, h_next, l.pitch, orow, 1, ld_out, batch,
                                     static_cast<int>(H), s), "vb_rnn_cell_forward");
      }
    }
  }

 protected:
  void SetExtra(const std::string& name, const Property& v) override {
    if (name == "hidden_size") hidden_ = static_cast<int64_t>(v.num);
    else if (name == "return_sequences") return_sequences_ = v.kind == Property::Bool ? v.b : v.num != 0;
    else if (name == "sequence_length") seq_len_ = static_cast<int64_t>(v.num);
  }

 private:
  struct Layout {
    int64_t pitch, zpitch;
    size_t xh, z, state, total;
  };
  static size_t Align(size_t v) { return (v + 255) / 256 * 256; }
  Layout MakeLayout(int batch, int64_t T, int64_t D) const {
    Layout l;
    l.pitch = (D + hidden_ + 7) / 8 * 8;
    l.zpitch = (GATES * hidden_ + 3) / 4 * 4;
    l.xh = 0;
    l.z = Align(static_cast<size_t>(T) * batch * l.pitch * 2);
    l.state = l.z + Align(static_cast<size_t>(batch) * l.zpitch * 4);
    const size_t state_floats = GATES == 4 ? static_cast<size_t>(batch) * hidden_ * 6 : static_cast<size_t>(batch) * hidden_;
    l.total = l.state + Align(state_floats * 4);
    return l;
  }
  void Geometry(const Shape& in, int64_t* T, int64_t* D) const {
    if (hidden_ <= 0) throw std::runtime_error(std::string(Name()) + ": hidden_size was not set");
    const int64_t n = shape_size(in);
    if (in.size() == 2 && (seq_len_ <= 0 || seq_len_ == in[0])) {
      *T = in[0];
      *D = in[1];
      return;
    }
    if (seq_len_ <= 0 || n % seq_len_) throw std::runtime_error(std::string(Name()) + ": input is not [T, D] and sequence_length does not divide it");
    *T = seq_len_;
    *D = n / seq_len_;
  }
  int64_t hidden_ = 0, seq_len_ = 0;
  bool return_sequences_ = true;
};

class RNN : public RecurrentUnit<1> {
 public:
  const char* Uuid() const override { return "0d2a3e57-7b0f-4c63-9b55-5a4e3d6f2c11"; }
  const char* Name() const override { return "RNN"; }
};
VELES_REGISTER_UNIT(RNN, "0d2a3e57-7b0f-4c63-9b55-5a4e3d6f2c11", "RNN");

class LSTM : public RecurrentUnit<4> {
 public:
  const char* Uuid() const override { return "6c1e8f0a-2d47-4b39-8a6e-9f3b7c5d1e24"; }
  const char* Name() const override { return "LSTM"; }
};
VELES_REGISTER_UNIT(LSTM, "6c1e8f0a-2d47-4b39-8a6e-9f3b7c5d1e24", "LSTM");

// ---------------------------------------------------------------------------------------------------------
class LRNormalizerForward : public Unit {
 public:
  const char* Uuid() const override { return "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0021"; }
  const char* Name() const override { return "LRNormalizerForward"; }
  void SetParameter(const std::string& name, const Property& v) override {
    if (name == "alpha") alpha_ = static_cast<float>(v.num);
    else if (name == "beta") beta_ = static_cast<float>(v.num);
    else if (name == "k") k_ = static_cast<float>(v.num);
    else if (name == "n") n_ = static_cast<int>(v.num);
  }
  Shape OutputShape(const Shape& in) const override { return in; }
  void Execute(const void* in, void* out, void*, int batch, const Shape& s, cudaStream_t st) override {
    const int C = static_cast<int>(s.back());
    vb_check(vb_lrn_forward(in, out, 1, static_cast<long long>(batch) * shape_size(s) / C, C, n_, alpha_, beta_, k_, st),
             "vb_lrn_forward");
  }

 private:
  float alpha_ = 1e-4f, beta_ = 0.75f, k_ = 2.0f;
  int n_ = 5;
};
VELES_REGISTER_UNIT(LRNormalizerForward, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0021", "LRNormalizerForward");

// Dropout is the identity at inference time.
class DropoutForward : public Unit {
 public:
  const char* Uuid() const override { return "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0022"; }
  const char* Name() const override { return "DropoutForward"; }
  void SetParameter(const std::string&, const Property&) override {}
  Shape OutputShape(const Shape& in) const override { return in; }
  void Execute(const void* in, void* out, void*, int batch, const Shape& s, cudaStream_t st) override {
    VU_CUDA(cudaMemcpyAsync(out, in, static_cast<size_t>(batch) * shape_size(s) * 2, cudaMemcpyDeviceToDevice, st));
  }
};
VELES_REGISTER_UNIT(DropoutForward, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0022", "DropoutForward");

// Stand-alone activation (activation code + factor come from the package).
class ActivationForward : public Unit {
 public:
  const char* Uuid() const override { return "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0023"; }
  const char* Name() const override { return "ActivationForward"; }
  void SetParameter(const std::string& name, const Property& v) override {
    if (name == "activation") code_ = static_cast<int>(v.num);
    else if (name == "factor") factor_ = static_cast<float>(v.num);
  }
  Shape OutputShape(const Shape& in) const override { return in; }
  void Execute(const void* in, void* out, void*, int batch, const Shape& s, cudaStream_t st) override {
    vb_check(vb_activation_forward(code_, in, out, static_cast<long long>(batch) * shape_size(s), 1, factor_, st),
             "vb_activation_forward");
  }

 private:
  int code_ = 0;
  float factor_ = 1.0f;
};
VELES_REGISTER_UNIT(ActivationForward, "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0023", "ActivationForward");

}  // namespace veles
