// LADSPA-shaped streaming host over the C ABI (include/bsrnn_b200_plugin.h).
// Behaviour follows speech-ladspa-onnx.cpp:152-267 (FIFO, one-hop delay, input L only, dry/wet
// control, state carry) and :271-342 (descriptor); the signal work is bsb_stream_step_mix, replayed
// as one CUDA graph per hop (H2D of the chunk + the mix value, the step's kernels, D2H of 1024 samples).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/bsrnn_b200.h"
#include "../../include/bsrnn_b200_plugin.h"

namespace {

constexpr int kRows = 2;                                  // n_channels (speech-ladspa-onnx.cpp:14)
constexpr int kChunk = BSB_HOP;                           // chunk_size (:9)
constexpr long long kStateFloats = 4LL * 2 * kRows * BSB_N_BANDS * BSB_FEAT;

#define PCK(expr)                                 \
  do {                                            \
    cudaError_t e_ = (expr);                      \
    if (e_ != cudaSuccess) return (int)e_;        \
  } while (0)
#define PCI(expr)                                 \
  do {                                            \
    int e_ = (expr);                              \
    if (e_ != 0) return e_;                       \
  } while (0)

long long weight_floats() {
  long long n = 0;
  for (int i = 0; i < BSB_N_TENSORS; ++i) n += bsb_tensor_numel(i);
  return n;
}

}  // namespace

struct bsb_plugin {
  int device = 0;
  cudaStream_t stream = nullptr;
  // device memory
  void* packed = nullptr;
  void* ws = nullptr;
  size_t ws_bytes = 0;
  float* d_in = nullptr;        // [2][1024] chunk rows + 1 float mix
  float* d_ring = nullptr;      // [2][2048]
  float* d_tail = nullptr;      // [2][1024]
  float* d_out = nullptr;       // [2][1024]
  float* d_state[2] = {nullptr, nullptr};
  // pinned host staging
  float* h_in = nullptr;        // same shape as d_in
  float* h_out = nullptr;       // [1024] (row 0: the reference writes one signal to both outputs)
  // FIFO (ladspa_buffer_in/out, ladspa_pos: speech-ladspa-onnx.cpp:48-50)
  float fifo_in[kChunk];
  float fifo_out[kChunk];
  int pos = 0;
  // ports
  float* port[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  // per-hop graphs (the state ping-pongs between two buffers -> one graph per parity)
  cudaGraphExec_t graph[2] = {nullptr, nullptr};
  long long hops = 0;
  double last_ms = 0.0, max_ms = 0.0;

  int enqueue(int parity) {
    const size_t in_bytes = (size_t)(kRows * kChunk + 1) * sizeof(float);
    PCK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, stream));
    PCI(bsb_stream_step_mix(d_in, kRows, nullptr, packed, d_ring, d_state[parity], d_state[parity ^ 1], d_tail,
                            d_out, d_in + kRows * kChunk, ws, ws_bytes, stream));
    PCK(cudaMemcpyAsync(h_out, d_out, kChunk * sizeof(float), cudaMemcpyDeviceToHost, stream));
    return 0;
  }

  int capture(int parity) {
    cudaGraph_t g = nullptr;
    PCK(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
    const int e = enqueue(parity);
    const cudaError_t ce = cudaStreamEndCapture(stream, &g);
    if (e != 0) { if (g) cudaGraphDestroy(g); return e; }
    PCK(ce);
    const cudaError_t ie = cudaGraphInstantiate(&graph[parity], g, 0);
    cudaGraphDestroy(g);
    PCK(ie);
    return 0;
  }

  // new_chunk(), speech-ladspa-onnx.cpp:193-267
  int hop() {
    const auto t0 = std::chrono::steady_clock::now();
    PCK(cudaSetDevice(device));
    const float mix = port[BSB_PORT_CONTROL] ? *port[BSB_PORT_CONTROL] : 1.0f;
    // both model rows carry input L (:199-203: every channel of current_chunk is filled from buffer 0)
    std::memcpy(h_in, fifo_in, sizeof(fifo_in));
    std::memcpy(h_in + kChunk, fifo_in, sizeof(fifo_in));
    h_in[kRows * kChunk] = mix;
    const int parity = (int)(hops & 1);
    if (hops < 2) {                 // first use of each parity runs eagerly (lazy table uploads are not capturable)
      PCI(enqueue(parity));
    } else {
      if (!graph[parity]) PCI(capture(parity));
      PCK(cudaGraphLaunch(graph[parity], stream));
    }
    PCK(cudaStreamSynchronize(stream));
    std::memcpy(fifo_out, h_out, sizeof(fifo_out));
    pos = 0;
    ++hops;
    last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (last_ms > max_ms) max_ms = last_ms;
    return 0;
  }

  int reset() {
    PCK(cudaSetDevice(device));
    PCK(cudaMemsetAsync(d_ring, 0, sizeof(float) * kRows * BSB_N_FFT, stream));
    PCK(cudaMemsetAsync(d_tail, 0, sizeof(float) * kRows * kChunk, stream));
    PCK(cudaMemsetAsync(d_state[0], 0, sizeof(float) * kStateFloats, stream));
    PCK(cudaMemsetAsync(d_state[1], 0, sizeof(float) * kStateFloats, stream));
    PCK(cudaStreamSynchronize(stream));
    std::memset(fifo_in, 0, sizeof(fifo_in));
    std::memset(fifo_out, 0, sizeof(fifo_out));
    pos = 0;
    // keep hop parity consistent with the graphs: the live state must sit in d_state[hops & 1]; both are zero now
    last_ms = max_ms = 0.0;
    return 0;
  }

  int init(const float* weights, long long n_floats, int dev) {
    if (!weights || n_floats != weight_floats() || dev < 0) return BSB_E_ARG;
    device = dev;
    PCK(cudaSetDevice(device));
    PCI(bsb_init(device));           // window / twiddle tables: nothing allocates on the audio thread later
    PCK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    // flat image -> 288 device pointers -> packed buffer
    float* d_flat = nullptr;
    PCK(cudaMalloc(&d_flat, sizeof(float) * n_floats));
    PCK(cudaMemcpyAsync(d_flat, weights, sizeof(float) * n_floats, cudaMemcpyHostToDevice, stream));
    std::vector<const void*> table(BSB_N_TENSORS, nullptr);
    long long off = 0;
    for (int i = 0; i < BSB_N_TENSORS; ++i) {
      const long long n = bsb_tensor_numel(i);
      table[i] = n > 0 ? d_flat + off : nullptr;
      off += n;
    }
    PCK(cudaMalloc(&packed, bsb_packed_weights_bytes()));
    int e = bsb_pack_weights(table.data(), BSB_N_TENSORS, packed, stream);
    cudaError_t se = cudaStreamSynchronize(stream);
    cudaFree(d_flat);
    PCI(e);
    PCK(se);
    ws_bytes = bsb_stream_workspace_bytes(kRows);
    PCK(cudaMalloc(&ws, ws_bytes));
    PCK(cudaMalloc(&d_in, sizeof(float) * (kRows * kChunk + 4)));
    PCK(cudaMalloc(&d_ring, sizeof(float) * kRows * BSB_N_FFT));
    PCK(cudaMalloc(&d_tail, sizeof(float) * kRows * kChunk));
    PCK(cudaMalloc(&d_out, sizeof(float) * kRows * kChunk));
    PCK(cudaMalloc(&d_state[0], sizeof(float) * kStateFloats));
    PCK(cudaMalloc(&d_state[1], sizeof(float) * kStateFloats));
    PCK(cudaMallocHost(&h_in, sizeof(float) * (kRows * kChunk + 4)));
    PCK(cudaMallocHost(&h_out, sizeof(float) * kChunk));
    return reset();
  }

  ~bsb_plugin() {
    cudaSetDevice(device);
    for (int i = 0; i < 2; ++i) if (graph[i]) cudaGraphExecDestroy(graph[i]);
    if (stream) cudaStreamSynchronize(stream);
    cudaFree(packed); cudaFree(ws); cudaFree(d_in); cudaFree(d_ring); cudaFree(d_tail); cudaFree(d_out);
    cudaFree(d_state[0]); cudaFree(d_state[1]);
    if (h_in) cudaFreeHost(h_in);
    if (h_out) cudaFreeHost(h_out);
    if (stream) cudaStreamDestroy(stream);
  }
};

extern "C" {

long long bsb_plugin_weight_floats(void) { return weight_floats(); }

int bsb_plugin_create(bsb_plugin** out, const float* weights, long long n_floats, int device) {
  if (!out) return BSB_E_ARG;
  *out = nullptr;
  bsb_plugin* p = new (std::nothrow) bsb_plugin();
  if (!p) return (int)cudaErrorMemoryAllocation;
  const int e = p->init(weights, n_floats, device);
  if (e != 0) { delete p; return e; }
  *out = p;
  return 0;
}

int bsb_plugin_create_from_file(bsb_plugin** out, const char* path, int device) {
  if (!out || !path) return BSB_E_ARG;
  *out = nullptr;
  const long long n = weight_floats();
  FILE* f = std::fopen(path, "rb");
  if (!f) return BSB_E_ARG;
  std::vector<float> w((size_t)n);
  const size_t got = std::fread(w.data(), sizeof(float), (size_t)n, f);
  const bool exact = got == (size_t)n && std::fgetc(f) == EOF;
  std::fclose(f);
  if (!exact) return BSB_E_SHAPE;
  return bsb_plugin_create(out, w.data(), n, device);
}

void bsb_plugin_connect_port(bsb_plugin* p, unsigned long port, float* data) {
  if (p && port < 5) p->port[port] = data;
}

int bsb_plugin_run(bsb_plugin* p, unsigned long sample_count) {
  if (!p) return BSB_E_ARG;
  const float* in_l = p->port[BSB_PORT_IN_L];
  float* out_l = p->port[BSB_PORT_OUT_L];
  float* out_r = p->port[BSB_PORT_OUT_R];
  if (sample_count > 0 && !in_l) return BSB_E_ARG;
  unsigned long done = 0;
  while (sample_count > 0) {
    const unsigned long room = (unsigned long)(kChunk - p->pos);
    const unsigned long n = sample_count < room ? sample_count : room;
    std::memcpy(p->fifo_in + p->pos, in_l + done, n * sizeof(float));
    if (out_l) std::memcpy(out_l + done, p->fifo_out + p->pos, n * sizeof(float));
    if (out_r) std::memcpy(out_r + done, p->fifo_out + p->pos, n * sizeof(float));
    p->pos += (int)n;
    done += n;
    sample_count -= n;
    if (p->pos == kChunk) PCI(p->hop());
  }
  return 0;
}

int bsb_plugin_reset(bsb_plugin* p) { return p ? p->reset() : BSB_E_ARG; }
long long bsb_plugin_hops(const bsb_plugin* p) { return p ? p->hops : -1; }
double bsb_plugin_last_hop_ms(const bsb_plugin* p) { return p ? p->last_ms : -1.0; }
double bsb_plugin_max_hop_ms(const bsb_plugin* p) { return p ? p->max_ms : -1.0; }

int bsb_plugin_get_state(bsb_plugin* p, float* state_host, long long n_floats) {
  if (!p || !state_host || n_floats != kStateFloats) return BSB_E_ARG;
  PCK(cudaSetDevice(p->device));
  PCK(cudaMemcpy(state_host, p->d_state[p->hops & 1], sizeof(float) * kStateFloats, cudaMemcpyDeviceToHost));
  return 0;
}

void bsb_plugin_destroy(bsb_plugin* p) { delete p; }

}  // extern "C"

// ------------------------------------------------------------------------------------------
// LADSPA glue (speech-ladspa-onnx.cpp:271-342)
namespace {

bsb_ladspa_handle ld_instantiate(const bsb_ladspa_descriptor*, unsigned long /*sample_rate: ignored, :271*/) {
  const char* path = std::getenv("BSRNN_B200_WEIGHTS");
  const char* dev = std::getenv("BSRNN_B200_DEVICE");
  if (!path) {
    std::fprintf(stderr, "bsrnn_b200 plugin: BSRNN_B200_WEIGHTS is not set\n");
    return nullptr;
  }
  bsb_plugin* p = nullptr;
  const int e = bsb_plugin_create_from_file(&p, path, dev ? std::atoi(dev) : 0);
  if (e != 0) {
    std::fprintf(stderr, "bsrnn_b200 plugin: cannot start from %s: %s\n", path, bsb_error_string(e));
    return nullptr;
  }
  return p;
}
void ld_connect(bsb_ladspa_handle h, unsigned long port, bsb_ladspa_data* data) {
  bsb_plugin_connect_port(static_cast<bsb_plugin*>(h), port, data);
}
void ld_run(bsb_ladspa_handle h, unsigned long n) {
  const int e = bsb_plugin_run(static_cast<bsb_plugin*>(h), n);
  if (e != 0) std::fprintf(stderr, "bsrnn_b200 plugin: run failed: %s\n", bsb_error_string(e));
}
void ld_cleanup(bsb_ladspa_handle h) { bsb_plugin_destroy(static_cast<bsb_plugin*>(h)); }

const int kPortDesc[5] = {
    BSB_LADSPA_PORT_INPUT | BSB_LADSPA_PORT_CONTROL, BSB_LADSPA_PORT_INPUT | BSB_LADSPA_PORT_AUDIO,
    BSB_LADSPA_PORT_INPUT | BSB_LADSPA_PORT_AUDIO,   BSB_LADSPA_PORT_OUTPUT | BSB_LADSPA_PORT_AUDIO,
    BSB_LADSPA_PORT_OUTPUT | BSB_LADSPA_PORT_AUDIO};
const char* const kPortNames[5] = {"Control", "Input (Left)", "Input (Right)", "Output (Left)", "Output (Right)"};
const bsb_ladspa_port_range_hint kHints[5] = {
    {BSB_LADSPA_HINT_DEFAULT_1 | BSB_LADSPA_HINT_BOUNDED_BELOW, 0.f, 0.f}, {0, 0.f, 0.f}, {0, 0.f, 0.f},
    {0, 0.f, 0.f}, {0, 0.f, 0.f}};
const bsb_ladspa_descriptor kDescriptor = {
    0xf433b044UL, "speech_separator", 0, "Speech Separator (B200)", "speechseparation_b200", "None", 5,
    kPortDesc, kPortNames, kHints, nullptr,
    ld_instantiate, ld_connect, nullptr, ld_run, nullptr, nullptr, nullptr, ld_cleanup};

}  // namespace

extern "C" const bsb_ladspa_descriptor* ladspa_descriptor(unsigned long idx) {
  return idx == 0 ? &kDescriptor : nullptr;
}
