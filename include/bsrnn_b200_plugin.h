/*
 * bsrnn_b200_plugin.h -- the LADSPA-shaped streaming host of libbsrnn_b200.so.
 *
 * Replaces class SpeechSeparator + the LADSPA glue of the reference plugin
 * (speech-ladspa-onnx.cpp:17-267 and :271-342): the same 1024-sample FIFO, the same port
 * numbering and dry/wet control, the same one-hop output delay -- but FFTW + Ort::Session::Run
 * (speech-ladspa-onnx.cpp:198-211,236) become ONE CUDA graph per hop around bsb_stream_step_mix.
 *
 * ladspa.h is not installed in this image, so the descriptor type is declared here with the
 * field order and types of the LADSPA 1.1 struct (what a host dlopen()ing the library expects).
 *
 * All pointers in THIS header are HOST pointers (the plugin owns its device memory).
 * Thread model as in the reference: one audio thread calls run(); no internal threads.
 * Errors never cross as exceptions: int returns (0 ok, <0 BSB_E_*, >0 cudaError_t); the LADSPA
 * callbacks themselves report to stderr only, like the reference (speech-ladspa-onnx.cpp:173).
 */
#ifndef BSRNN_B200_PLUGIN_H
#define BSRNN_B200_PLUGIN_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef float bsb_ladspa_data;              /* LADSPA_Data */
typedef void* bsb_ladspa_handle;            /* LADSPA_Handle */

typedef struct bsb_ladspa_port_range_hint {  /* LADSPA_PortRangeHint */
  int   HintDescriptor;
  float LowerBound;
  float UpperBound;
} bsb_ladspa_port_range_hint;

typedef struct bsb_ladspa_descriptor {       /* LADSPA_Descriptor, same field order */
  unsigned long UniqueID;
  const char*   Label;
  int           Properties;
  const char*   Name;
  const char*   Maker;
  const char*   Copyright;
  unsigned long PortCount;
  const int*    PortDescriptors;
  const char* const* PortNames;
  const bsb_ladspa_port_range_hint* PortRangeHints;
  void*         ImplementationData;
  bsb_ladspa_handle (*instantiate)(const struct bsb_ladspa_descriptor*, unsigned long sample_rate);
  void (*connect_port)(bsb_ladspa_handle, unsigned long port, bsb_ladspa_data* data);
  void (*activate)(bsb_ladspa_handle);
  void (*run)(bsb_ladspa_handle, unsigned long sample_count);
  void (*run_adding)(bsb_ladspa_handle, unsigned long sample_count);
  void (*set_run_adding_gain)(bsb_ladspa_handle, bsb_ladspa_data gain);
  void (*deactivate)(bsb_ladspa_handle);
  void (*cleanup)(bsb_ladspa_handle);
} bsb_ladspa_descriptor;

/* LADSPA constants used by the descriptor (values of ladspa.h 1.1) */
#define BSB_LADSPA_PORT_INPUT    0x1
#define BSB_LADSPA_PORT_OUTPUT   0x2
#define BSB_LADSPA_PORT_CONTROL  0x4
#define BSB_LADSPA_PORT_AUDIO    0x8
#define BSB_LADSPA_HINT_BOUNDED_BELOW 0x1
#define BSB_LADSPA_HINT_DEFAULT_1     0x240

/* port numbers (speech-ladspa-onnx.cpp:152-172) */
#define BSB_PORT_CONTROL 0
#define BSB_PORT_IN_L    1
#define BSB_PORT_IN_R    2
#define BSB_PORT_OUT_L   3
#define BSB_PORT_OUT_R   4

/* The LADSPA entry point (speech-ladspa-onnx.cpp:339-342): idx 0 -> the descriptor, else NULL.
 * instantiate() through the descriptor reads the weights from the file named by the environment
 * variable BSRNN_B200_WEIGHTS (the reference hard-codes a model.onnx path, :72) on CUDA device
 * BSRNN_B200_DEVICE (default 0); returns NULL and writes to stderr on failure. */
const bsb_ladspa_descriptor* ladspa_descriptor(unsigned long idx);

/* --- the same object without the descriptor indirection (what the tests and bench drive) --- */
typedef struct bsb_plugin bsb_plugin;

/* number of floats of the flat weight image: the 288 state_dict tensors in reference order,
 * each C-contiguous, back to back (7,481,062). */
long long bsb_plugin_weight_floats(void);
/* weights: HOST float[n_floats] flat image. device: CUDA ordinal. */
int  bsb_plugin_create(bsb_plugin** out, const float* weights, long long n_floats, int device);
/* raw little-endian fp32 file of exactly bsb_plugin_weight_floats() floats */
int  bsb_plugin_create_from_file(bsb_plugin** out, const char* path, int device);
/* SpeechSeparator::connect (speech-ladspa-onnx.cpp:152-172); unknown ports are ignored */
void bsb_plugin_connect_port(bsb_plugin* p, unsigned long port, float* data);
/* SpeechSeparator::run (:174-191): any sample_count; every 1024 collected samples one hop runs
 * (:193-267). Only input L is analysed and the result goes to both outputs, as in the reference.
 * A NULL control port means mix = 1. Returns the first error of a hop (state is then undefined). */
int  bsb_plugin_run(bsb_plugin* p, unsigned long sample_count);
/* zero ring/state/FIFO: back to the state after create */
int  bsb_plugin_reset(bsb_plugin* p);
/* hops executed, and host-visible wall time of the last / the slowest hop in milliseconds
 * (H2D of 1024 samples + graph launch + D2H + sync) */
long long bsb_plugin_hops(const bsb_plugin* p);
double    bsb_plugin_last_hop_ms(const bsb_plugin* p);
double    bsb_plugin_max_hop_ms(const bsb_plugin* p);
/* copies the carried LSTM state [4][2][24][64] to host (parity tests) */
int  bsb_plugin_get_state(bsb_plugin* p, float* state_host, long long n_floats);
void bsb_plugin_destroy(bsb_plugin* p);

#ifdef __cplusplus
}
#endif
#endif /* BSRNN_B200_PLUGIN_H */
