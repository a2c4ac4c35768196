/*
 * honk_b200.h - C ABI of libhonk_b200.so, the B200 (sm_100a) implementation of Honk's
 * batched keyword-spotting inference path.
 *
 * The reference (ljj7975/honk) is pure Python and has no FFI; this header IS the boundary
 * a maintainer binds (ctypes stub in INTEGRATION.md).  Each entry point names the
 * reference interface it replaces (paths relative to the reference tree).
 *
 * Conventions
 *   - every function returns 0 on success or a negative honk_status; the message for the
 *     calling thread's last failure is honk_last_error().  Nothing throws across the ABI.
 *   - pointers named d_* are DEVICE pointers owned by the caller (PyTorch tensors in the
 *     Python shim); h_* are HOST pointers.  The library owns only model handles.
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it and nothing
 *     synchronises, so calls are re-entrant across streams/threads (a model handle is
 *     immutable after creation; scratch memory is the caller-provided workspace).
 *   - there is NO CPU fallback: a missing/unsupported GPU is an error (HONK_ERR_DEVICE).
 */
#ifndef HONK_B200_H_
#define HONK_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HONK_B200_ABI_VERSION 2

#if defined(__GNUC__)
#define HONK_API __attribute__((visibility("default")))
#else
#define HONK_API
#endif

typedef enum honk_status {
  HONK_OK = 0,
  HONK_ERR_INVALID = -1,   /* bad argument (shape, dtype, null pointer)            */
  HONK_ERR_KEY = -2,       /* state_dict key missing / unexpected / wrong shape    */
  HONK_ERR_DEVICE = -3,    /* no CUDA device or not sm_100                         */
  HONK_ERR_CUDA = -4,      /* a CUDA runtime/driver call failed                    */
  HONK_ERR_WORKSPACE = -5, /* caller workspace too small                           */
  HONK_ERR_UNSUPPORTED = -6
} honk_status;

typedef enum honk_pcm_dtype {
  HONK_PCM_I16 = 0,        /* int16 wire format, scaled by 1/32768 (service.py:98) */
  HONK_PCM_F32 = 1,        /* float32 in [-1,1]  (model.py:295-303)                */
  HONK_PCM_F64 = 2         /* float64 in [-1,1]  (service.py:98 result)            */
} honk_pcm_dtype;

HONK_API int honk_abi_version(void);
HONK_API const char* honk_last_error(void);

/* Verifies that `device` is a CUDA device of compute capability 10.x. */
HONK_API int honk_device_check(int device);

/* ------------------------------------------------------------------------------------------
 * Front end.  Replaces AudioPreprocessor.__init__ / compute_mfccs
 * (utils/manage_audio.py:19-28, 30-42) and the per-clip loop of SpeechDataset.collate_fn
 * (utils/model.py:253-265).
 * ---------------------------------------------------------------------------------------- */
typedef struct honk_frontend_config {
  int32_t sr;              /* 16000 */
  int32_t n_fft;           /* 480   (multiple of 4, <= 512)        */
  int32_t hop_length;      /* 160   = sr / 1000 * hop_ms           */
  int32_t n_mels;          /* 40    (<= 64)                        */
  int32_t n_dct_filters;   /* 40    (<= 64)                        */
  float f_min;             /* 20    */
  float f_max;             /* 4000  */
} honk_frontend_config;

typedef struct honk_frontend honk_frontend;

/* Builds the window / DFT / mel / DCT tables in float64 on the host exactly as
 * librosa 0.6 does and uploads them to `device`. */
HONK_API int honk_frontend_create(const honk_frontend_config* cfg, int device, honk_frontend** out);
HONK_API int honk_frontend_destroy(honk_frontend* fe);
/* frames produced for n_samples input samples: 1 + n_samples / hop_length (centre=True) */
HONK_API int64_t honk_frontend_num_frames(const honk_frontend* fe, int64_t n_samples);
/* host copies of the tables: dct [n_dct, n_mels], mel [n_mels, 1 + n_fft/2], float64 */
HONK_API int honk_frontend_get_dct(const honk_frontend* fe, double* h_out);
HONK_API int honk_frontend_get_mel(const honk_frontend* fe, double* h_out);

/* d_pcm: [n_clips, clip_stride] elements of `dtype`, first n_samples of each row used; clip_stride <
 * n_samples is allowed (overlapping rows = sliding windows over one buffer, service.py:106-110).
 * d_mfcc: float32 [n_clips, n_frames, n_dct_filters] (time-major, the layout
 * collate_fn's reshape(1,-1,40) and service.py's squeeze(2) give the model). */
HONK_API int honk_frontend_mfcc(const honk_frontend* fe, const void* d_pcm, int dtype, int64_t n_clips,
                       int64_t n_samples, int64_t clip_stride, float* d_mfcc, void* d_workspace,
                       int64_t workspace_bytes, void* stream);
/* DEVICE scratch honk_frontend_mfcc needs for this call (256-byte aligned; 0 for the paths that need none).
 * int16 PCM with the default geometry goes through the integer-limb tensor-core DFT, which stages the float32
 * power spectrum (480 B per frame, at most 4096 clips at a time) between its two kernels. */
HONK_API int64_t honk_frontend_workspace_bytes(const honk_frontend* fe, int64_t n_clips, int64_t n_samples, int dtype);
/* Allocation-free variant of honk_frontend_mfcc for int16 PCM.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  It takes no workspace and calls no cudaMalloc / cudaFree or other allocating routine: the float32
 * power spectrum between its two kernels lives in a file-scope static device array of HONK_NOALLOC_MAX_FRAMES x 480 B
 * (3.1 MB, held by every GPU that loads the library until exit).  n_clips x frames per clip must not exceed
 * HONK_NOALLOC_MAX_FRAMES (64 one-second clips) or the call fails with HONK_ERR_INVALID; the front end must have the default
 * geometry (HONK_ERR_UNSUPPORTED otherwise).  Inputs and outputs as for honk_frontend_mfcc with dtype HONK_PCM_I16.  All
 * calls on a device share the array: calls that overlap on the GPU (two streams, two threads) race - serialise them. */
#define HONK_NOALLOC_MAX_FRAMES 6464
HONK_API int honk_frontend_mfcc_noalloc(const honk_frontend* fe, const int16_t* d_pcm, int64_t n_clips, int64_t n_samples,
                               int64_t clip_stride, float* d_mfcc, void* stream);
/* Name of the path int16 PCM takes through this front end: "tcgen05-i8-limbs" (exact integer DFT on the tensor
 * cores), "fp64-dmma" or "fp64-dfma". */
HONK_API const char* honk_frontend_path(const honk_frontend* fe);

/* ------------------------------------------------------------------------------------------
 * Models.  Replaces SpeechResModel / SpeechModel construction, load_state_dict and forward
 * (utils/model.py:72-80, 82-121, 123-205) for every entry of _configs (model.py:381-414).
 * ---------------------------------------------------------------------------------------- */
typedef enum honk_family { HONK_FAMILY_RES = 0, HONK_FAMILY_CNN = 1 } honk_family;

/* Plain-integer mirror of a find_config() dict (model.py:59-62, 381-414).  Absent optional
 * keys are 0. */
typedef struct honk_model_config {
  int32_t family;            /* honk_family: name.startswith("res") (model.py:51-57)        */
  int32_t n_labels;
  int32_t height, width;     /* input frames x coefficients: 101 x 40                       */
  /* res* (model.py:85-102) */
  int32_t n_feature_maps;
  int32_t n_layers;
  int32_t use_dilation;
  int32_t res_pool[2];       /* {0,0} when the config has no "res_pool"                     */
  /* cnn-* (model.py:126-184) */
  int32_t n_feature_maps1;
  int32_t conv1_size[2], conv1_pool[2], conv1_stride[2];
  int32_t has_conv2;
  int32_t n_feature_maps2;
  int32_t conv2_size[2], conv2_pool[2], conv2_stride[2];
  int32_t tf_variant;        /* 0 => "lin" layer (-> 32) present, dnn1 followed by ReLU      */
  int32_t dnn1_size, dnn2_size;  /* 0 when absent                                           */
  /* build options (extensions, default 0) */
  int32_t precision;         /* 0 = f16 tensor-core operands + f32 accumulate where a tcgen05
                                path exists; 1 = force the all-f32 CUDA-core path           */
} honk_model_config;

/* One state_dict entry under the reference's key name (model.py:85-102). */
typedef struct honk_tensor {
  const char* name;          /* e.g. "conv3.weight", "bn3.running_var", "output.bias"       */
  const void* h_data;        /* host, contiguous, float32 (int64 for num_batches_tracked)   */
  int32_t ndim;
  int64_t dims[4];
} honk_tensor;

typedef struct honk_model honk_model;

/* Strict load (like load_state_dict, model.py:79-80): every expected key must be present
 * with the right shape and no unknown key may appear (num_batches_tracked is accepted and
 * ignored).  BatchNorm statistics are folded and weights repacked for the kernels here. */
HONK_API int honk_model_create(const honk_model_config* cfg, const honk_tensor* tensors, int32_t n_tensors,
                      int device, honk_model** out);
HONK_API int honk_model_destroy(honk_model* m);

/* Scratch bytes honk_model_forward* needs for a batch of n_clips (256-byte aligned). */
HONK_API int64_t honk_model_workspace_bytes(const honk_model* m, int64_t n_clips);

/* d_mfcc float32 [n_clips, height, width] -> d_logits float32 [n_clips, n_labels]
 * (pre-softmax, like forward(); model.py:104-121, 186-205).  `height` x `width` is the shape of one
 * clip of d_mfcc as the caller holds it; it must equal the model's input (101 x 40 by default) or the
 * call fails with HONK_ERR_INVALID - the reference raises a size mismatch for a wrong shape
 * (model.py:195-205), it never reads past the tensor. */
HONK_API int honk_model_forward(const honk_model* m, const float* d_mfcc, int64_t n_clips, int32_t height, int32_t width,
                       float* d_logits, void* d_workspace, int64_t workspace_bytes, void* stream);

/* Fused path: PCM -> front end -> CNN without materialising MFCCs for the caller.
 * Same semantics as honk_frontend_mfcc followed by honk_model_forward. */
HONK_API int honk_model_forward_pcm(const honk_model* m, const honk_frontend* fe, const void* d_pcm, int dtype,
                           int64_t n_clips, int64_t n_samples, int64_t clip_stride, float* d_logits,
                           void* d_workspace, int64_t workspace_bytes, void* stream);
HONK_API int64_t honk_model_workspace_bytes_pcm(const honk_model* m, const honk_frontend* fe, int64_t n_clips,
                                       int64_t n_samples);

/* Host-buffer pipeline, the call a serving process makes with audio that sits in host memory
 * (the role of SpeechDataset.collate_fn + model(...).cpu() in utils/train.py:41-46, 71-80 and of
 * TorchLabelService.label in service.py:88-104): h_pcm [n_clips, clip_stride] of `dtype` in HOST
 * memory -> h_logits float32 [n_clips, n_labels] in HOST memory.  The batch is cut into chunks
 * and the host->device copy of chunk i+1 runs on a second stream while chunk i is in the
 * kernels; page-locked h_pcm / h_logits make those copies truly asynchronous (pageable memory
 * works, without the overlap).  Returns after h_logits is complete.  d_workspace is DEVICE
 * scratch of honk_model_workspace_bytes_pcm_host bytes. */
HONK_API int honk_model_forward_pcm_host(const honk_model* m, const honk_frontend* fe, const void* h_pcm, int dtype,
                                int64_t n_clips, int64_t n_samples, int64_t clip_stride, float* h_logits,
                                void* d_workspace, int64_t workspace_bytes, void* stream);
/* TorchLabelService.label for a batch of requests in ONE call (service.py:88-104): host PCM in, host
 * (argmax label, softmax probability) out - copy, front end, model, softmax/argmax, copy back, one
 * stream synchronisation.  Same workspace as honk_model_forward_pcm_host. */
HONK_API int honk_model_label_pcm_host(const honk_model* m, const honk_frontend* fe, const void* h_pcm, int dtype,
                              int64_t n_clips, int64_t n_samples, int64_t clip_stride, int32_t* h_labels,
                              float* h_probs, void* d_workspace, int64_t workspace_bytes, void* stream);
HONK_API int64_t honk_model_workspace_bytes_pcm_host(const honk_model* m, const honk_frontend* fe, int64_t n_clips,
                                            int64_t n_samples, int dtype);

/* softmax + argmax + max-prob per clip: the tail of TorchLabelService.label
 * (service.py:103-104) and print_eval's argmax (utils/train.py:41-46).
 * d_labels int32 [n_clips], d_probs float32 [n_clips] (either may be NULL). */
HONK_API int honk_label_from_logits(const float* d_logits, int64_t n_clips, int32_t n_labels, int32_t* d_labels,
                           float* d_probs, void* stream);

/* Batched form of ListenEndpoint.POST (server.py:99-112) over service.py:106-110's stride():
 * d_pcm is ONE buffer of n_total_samples on the device; window w = samples [w*hop, w*hop + window)
 * (n = honk_listen_num_windows = 1 + (n_total - window) / hop, 0 when shorter than a window) is read
 * in place as a row of clip_stride = hop, so overlapping windows are never copied.  All windows go
 * through front end + model + softmax/argmax in one pass; then, in window order,
 *   d_sums[l]   float64 [n_labels]  sum of the winning probabilities of label l   (server.py:105-108)
 *   d_counts[l] int32   [n_labels]  windows won by label l (0 => the label is absent from the dict)
 *   d_first_hit int32   [1]         first window with label == keyword && prob >= min_keyword_prob,
 *                                   -1 if none (method "command_tagging", server.py:109-110). */
HONK_API int64_t honk_listen_num_windows(int64_t n_total_samples, int64_t window, int64_t hop);
HONK_API int64_t honk_listen_workspace_bytes(const honk_model* m, const honk_frontend* fe, int64_t n_total_samples,
                                    int64_t window, int64_t hop);
HONK_API int honk_listen(const honk_model* m, const honk_frontend* fe, const void* d_pcm, int dtype,
                int64_t n_total_samples, int64_t window, int64_t hop, int32_t keyword, float min_keyword_prob,
                double* d_sums, int32_t* d_counts, int32_t* d_first_hit, void* d_workspace,
                int64_t workspace_bytes, void* stream);

/* print_eval's accuracy numerator and nn.CrossEntropyLoss's sum for one batch of evaluate()
 * (utils/train.py:41-46, 66-83): d_correct[0] = #(argmax(logits_i) == labels_i) (first maximum, like
 * torch.max), d_loss_sum[0] = sum_i (logsumexp(logits_i) - logits_i[labels_i]) in float64; the caller
 * divides by n_clips.  d_labels int64 [n_clips] (torch.tensor(y), model.py:265); d_pred int32
 * [n_clips] may be NULL. */
HONK_API int honk_eval_metrics(const float* d_logits, const int64_t* d_labels, int64_t n_clips, int32_t n_labels,
                      int32_t* d_pred, int64_t* d_correct, double* d_loss_sum, void* stream);

/* The arithmetic of SpeechDataset.load_audio for a batch (utils/model.py:281-303), with the random
 * draws made by the caller: row i of d_pcm (float32, d_lengths[i] valid samples, NULL = all) is
 * zero-padded to input_length (:297), shifted by d_shifts[i] samples (_timeshift_audio, :267-273:
 * out[t] = x[t + shift], zeros outside; NULL = no shift), and where d_noise_gain[i] >= 0 mixed as
 * clip(gain * bg[offset_i + t] + x, -1, 1) (:299-303) in float32 with separately rounded multiply
 * and add, exactly like numpy; a negative gain leaves the clip unmixed and unclipped.  A silence
 * example (:293) is a row with d_lengths[i] = 0.  d_out float32 [n_clips, input_length]. */
HONK_API int honk_load_audio(const float* d_pcm, int64_t n_clips, int64_t clip_stride, const int32_t* d_lengths,
                    const int32_t* d_shifts, const float* d_noise_gain, const int64_t* d_noise_offset,
                    const float* d_bg_noise, int32_t input_length, float* d_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * Training step (replaces the body of the reference's
 * training loop, utils/train.py:127-139: model.train(), scores = model(model_in), CrossEntropyLoss,
 * loss.backward(), torch.optim.SGD.step()).  FP32 on the CUDA cores; BatchNorm2d(affine=False) in
 * training mode (batch statistics, running statistics moved by momentum 0.1; utils/model.py:98).
 * cnn-* family (SpeechModel, utils/model.py:123-205): conv + bias + ReLU -> Dropout -> MaxPool (x2), Linear layers with Dropout,
 * same loss and optimiser; the Dropout keep masks are drawn by the library or handed in (honk_trainer_set_dropout*).
 * ------------------------------------------------------------------------------------------ */
typedef struct honk_trainer honk_trainer;

/* The arguments of torch.optim.SGD as train.py:100 passes them (dampening 0). */
typedef struct honk_sgd {
  float lr, momentum, weight_decay;
  int32_t nesterov;
} honk_sgd;

/* Same strict state_dict check as honk_model_create; the trainer owns device copies of the parameters,
 * their gradients, the momentum buffers and the BatchNorm running statistics. */
HONK_API int honk_trainer_create(const honk_model_config* cfg, const honk_tensor* tensors, int32_t n_tensors, int device,
                                 honk_trainer** out);
HONK_API int honk_trainer_destroy(honk_trainer* t);
/* Scratch bytes of one step on n_clips clips (the activations of every layer are kept for the backward pass). */
HONK_API int64_t honk_trainer_workspace_bytes(const honk_trainer* t, int64_t n_clips);
/* One step on d_mfcc float32 [n_clips, height, width] with d_labels int64 [n_clips] (2 <= n_clips <= 65536).
 * d_scores float32 [n_clips, n_labels] receives the training-mode scores (what train.py:132 hands to print_eval),
 * d_loss one float32, the mean CrossEntropyLoss of the batch BEFORE the update (train.py:134); either may be NULL.
 * A label outside [0, n_labels) makes the loss NaN (the reference raises IndexError; the Python host checks first). */
HONK_API int honk_trainer_step(honk_trainer* t, const float* d_mfcc, const int64_t* d_labels, int64_t n_clips, int32_t height,
                               int32_t width, const honk_sgd* opt, float* d_scores, float* d_loss, void* d_workspace,
                               int64_t workspace_bytes, void* stream);
/* train.py:141-143 builds a NEW torch.optim.SGD when the learning-rate schedule advances, which drops the momentum
 * buffers; this does the same (the next step initialises them with its gradient). */
HONK_API int honk_trainer_reset_optimizer(honk_trainer* t);
/* cnn-* family only: nn.Dropout (utils/model.py:183; applied after conv1, conv2, dnn1 and dnn2, :186-205).  prob is
 * config["dropout_prob"] (default of a new handle: 0.5); the library draws its keep masks from a counter-based hash of
 * (seed, step, site, element) - a different stream from torch's, so a run is reproducible here but not bit-comparable with
 * the reference's.  For a bit-level comparison hand the masks in: honk_trainer_set_dropout_masks gives device pointers to
 * uint8 keep masks (1 = keep) by site: [0] conv1 [n, C1, H1, W1], [1] conv2 [n, C2, H2, W2], [2] dnn1 [n, dnn1_size],
 * [3] dnn2 [n, dnn2_size] (entries of absent layers are ignored, NULL or missing entries are drawn by the library); they
 * serve the NEXT step only. */
HONK_API int honk_trainer_set_dropout(honk_trainer* t, float prob, uint64_t seed);
HONK_API int honk_trainer_set_dropout_masks(honk_trainer* t, const unsigned char* const* d_masks, int32_t n_masks);
/* Steps taken by this handle (what every bn<i>.num_batches_tracked has advanced by). */
HONK_API int64_t honk_trainer_steps(const honk_trainer* t);
/* Copies one tensor to the host (synchronises the device): a state_dict key ("conv3.weight", "bn2.running_var",
 * "output.bias"), or "grad:<parameter>" / "momentum:<parameter>" for the last step's gradient / the SGD buffer. */
HONK_API int honk_trainer_read(const honk_trainer* t, const char* name, float* h_out, int64_t numel);

/* Number of kernels the library has launched in this process (for bench.py's gpu_launches). */
HONK_API int64_t honk_launch_count(void);

/* Optional per-span device timing (CUDA events on the launching stream) used by bench.py for
 * the roofline line.  Spans: "frontend", "res_conv0", "res_conv", "res_tail", "f32_model".
 * honk_profile_get synchronises the recorded events and returns the summed milliseconds and
 * the number of spans since the last reset. */
HONK_API int honk_profile_enable(int on);
HONK_API int honk_profile_reset(void);
HONK_API int honk_profile_get(const char* span, double* ms_total, int64_t* n_spans);

/* Name of the code path honk_model_forward takes: "tcgen05-f16" or "cuda-core-f32". */
HONK_API const char* honk_model_path(const honk_model* m);

#ifdef __cplusplus
}
#endif
#endif /* HONK_B200_H_ */
