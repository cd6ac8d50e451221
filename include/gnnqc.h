/* gnnqc.h - C ABI of libgnnqc.so: B200 (sm_100a) scoring / attribution / training of the
 * sensor-window anomaly models of timohouben/GNN-XAI-TimeSeries-QualityControl.
 *
 * The reference has no FFI layer: its boundary is the Keras Model object.  Each entry point
 * below names the reference call it stands in for (paths relative to the reference root).
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *  - Every function returns 0 on success or a negative GNNQC_ERR_* code; the message for the
 *    calling thread's last failure is gnnqc_last_error().
 *  - Pointers named *_dev are device pointers on the model's GPU, owned by the caller.
 *    Pointers named *_host are host pointers.  `stream` is a cudaStream_t passed as void*
 *    (NULL = the legacy default stream).  Device entry points only enqueue work; they do not
 *    synchronise the stream.
 *  - A handle is bound to one GPU and is not thread-safe; use one handle per GPU / rank.
 *  - There is no CPU fallback: device entry points fail with GNNQC_ERR_STATE on a handle that
 *    was opened with device < 0 (host-only inspection handle), and with GNNQC_ERR_CUDA when
 *    no usable GPU is present.
 */
#ifndef GNNQC_H_
#define GNNQC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GNNQC_ABI_VERSION 1

#define GNNQC_OK 0
#define GNNQC_ERR_ARG (-1)    /* bad argument / shape */
#define GNNQC_ERR_CKPT (-2)   /* unreadable or unsupported checkpoint */
#define GNNQC_ERR_CUDA (-3)   /* CUDA runtime error */
/* (-4 is unused: the library issues no collective itself - the training exchange steps are the caller's all-reduces
 * between gnnqc_train_moments_*, gnnqc_train_grads_* and gnnqc_train_apply, see "training" below) */
#define GNNQC_ERR_STATE (-5)  /* call not valid for this handle */

typedef struct gnnqc_model gnnqc_model;

/* Architecture + metadata of a loaded model.  Mirrors the attributes the reference reads
 * off the Keras model: model.model_info (libs/test_model.py:22-25, libs/visualize.py:160-163),
 * model_type / model_normalization (libs/create_model.py:163-165). */
typedef struct gnnqc_model_desc {
  int32_t has_gcn;        /* 1 = GCNClassifier, 0 = BaselineClassifier */
  int32_t soilnet;        /* 1 = ds_type 'soilnet', 0 = 'cml' */
  int32_t n_features;     /* 2 (cml) or 3 (soilnet) */
  int32_t n_lstm;         /* 7 for every shipped model */
  int32_t lstm_in[8];
  int32_t lstm_hid[8];
  int32_t dense_units;
  int32_t n_model_info;   /* 3 or 4 */
  int32_t model_info[4];  /* timestep_before, timestep_after, batch_size, freq */
  int32_t seq_len;        /* (timestep_before + timestep_after) / freq + 1 */
  int64_t n_params;       /* floats in the flat parameter vector */
  int64_t n_trainable;
  char ds_type[16];
  char normalization[32];
} gnnqc_model_desc;

int gnnqc_abi_version(void);
const char* gnnqc_last_error(void);

/* ---- model lifetime ------------------------------------------------------------------- */

/* Replaces tf.keras.models.load_model(model_path, compile=False) (notebooks/pipeline.ipynb:202,
 * :436; xai/libs/integrated_gradients.py:199-201).  `savedmodel_dir` is a Keras SavedModel
 * directory (model_cml, model_cml_baseline, model_soilnet, model_soilnet_baseline); only its
 * variables/variables.{index,data-00000-of-00001} are read.  device >= 0 uploads the weights to
 * that GPU; device < 0 opens a host-only handle (inspection, save). */
int gnnqc_model_load(const char* savedmodel_dir, int device, gnnqc_model** out);

/* Replaces the constructors GCNClassifier(model_config, preproc_config) /
 * BaselineClassifier(...) (libs/create_model.py:140-213, :261-341): builds a model of the
 * given architecture from a caller-supplied flat parameter vector (layout: see
 * gnnqc_model_segment).  model_info may be NULL. */
int gnnqc_model_create(int has_gcn, int soilnet, int filter_1_size, int n_stacks, int dense_units,
                       const float* params_host, int64_t n_params, const int32_t* model_info,
                       int n_model_info, const char* normalization, int device, gnnqc_model** out);

int gnnqc_model_describe(const gnnqc_model* m, gnnqc_model_desc* out);

/* Flat parameter vector (host copies).  Stands in for model.get_weights()/set_weights(). */
int gnnqc_model_get_params(const gnnqc_model* m, float* out_host, int64_t n);
int gnnqc_model_set_params(gnnqc_model* m, const float* in_host, int64_t n);

/* Segment `idx` of the flat vector: the variable's attribute path in the reference's object
 * graph (e.g. "time_layer/time1/cell/kernel"), its offset, shape and trainability.
 * Returns GNNQC_ERR_ARG once idx is past the last segment. */
int gnnqc_model_segment(const gnnqc_model* m, int idx, char path[128], int64_t* offset,
                        int64_t shape[4], int32_t* rank, int32_t* trainable);

/* Replaces model.save(model_config.model_path) (notebooks/pipeline.ipynb cell 14): writes
 * <dir>/variables/variables.{index,data-00000-of-00001} in TF tensor-bundle format with the
 * reference's checkpoint keys and object-graph paths. */
int gnnqc_model_save(const gnnqc_model* m, const char* dir);

void gnnqc_model_free(gnnqc_model* m);

/* ---- scoring -------------------------------------------------------------------------- */

/* Replaces BaselineClassifier.call on a CML batch (libs/create_model.py:343-377, driven by
 * model.predict at libs/test_model.py:29 with wrapper_batch_mapping_baseline_cml,
 * libs/preprocessing_functions.py:756-760), and the LSTM-stack + head half of every other
 * model.  x_dev: f32 [n_seq, T, C] row-major, C = the model's first LSTM input width
 * (2 for model_cml_baseline, 3 for model_soilnet_baseline after graph_reshape).
 * scores_dev: f32 [n_seq]. */
/* The SoilNet baseline straight from the reference's batch layout (BaselineClassifier.call with graph_reshape,
 * libs/create_model.py:242-258,343-347; wrapper_batch_mapping_baseline_soilnet, libs/preprocessing_functions.py:905-916):
 * features_dev f32 [sum N_b * T, C] rows in (window, t, sensor) order, node_off_dev i32 [B+1]; the reshape to
 * (window, sensor) sequences happens inside the tile builder.  scores_dev: f32 [sum N_b]. */
int gnnqc_score_rows(gnnqc_model* m, const float* features_dev, const int32_t* node_off_dev, int64_t B,
                     int32_t T, int64_t n_nodes_total, float* scores_dev, void* stream);
int gnnqc_score_seq(gnnqc_model* m, const float* x_dev, int64_t n_seq, int32_t T,
                    float* scores_dev, void* stream);

/* Replaces GCNClassifier.call for ds_type 'cml' (libs/create_model.py:215-223,233-240), i.e.
 * model.predict on batches from prepare_batch_cml / wrapper_batch_mapping_cml
 * (libs/preprocessing_functions.py:638-666, :744-746).
 *   features_dev  f32 [sum_b T*N_b, 2], rows in (window, timestep, node) order   (:640-641)
 *   anom_dev      f32 [B, T, 2]  the target sensor's own series                  (:642)
 *   edges_dev     i64 [E_total, 2] per-window edge list (row = receiving node, col = sending
 *                 node, both local to the window), NOT tiled over timesteps      (:221, :651-661)
 *   node_off_dev  i32 [B+1] prefix sums of N_b;  edge_off_dev i64 [B+1] prefix sums of E_b
 *   n_nodes_total = node_off[B] (host copy, so the call needs no device->host read)
 *   scores_dev    f32 [B]
 * Every window gets a score (the reference silently drops windows whose pooled features sum
 * to exactly 0, libs/create_model.py:30-36). */
int gnnqc_score_cml(gnnqc_model* m, const float* features_dev, const float* anom_dev,
                    const int64_t* edges_dev, const int32_t* node_off_dev,
                    const int64_t* edge_off_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                    float* scores_dev, void* stream);

/* ---- scoring from resident raw series (on-device window builder) ------------------------ */

/* Replaces, for sweeps over a whole network, the reference's materialise-then-score pipeline:
 * create_tfrecords_dataset (libs/preprocessing_functions.py:388-412: window slicing, windows
 * whose target has a NaN are skipped, neighbours with a NaN in the window are dropped),
 * parse_cml_tfrecord_fn (:620-622: subtract each sensor's rolling median taken at the window's
 * target time), prepare_batch_cml (:638-666) and GCNClassifier.call.  The raw network stays in HBM;
 * a window is named by (target sensor, time index), so ~8 bytes per window cross PCIe.
 *
 * gnnqc_series_create copies to `device`:
 *   raw_host, rmed_host  f32 [n_sensors, n_times, 2]  TL_1/TL_2 and their rolling medians
 *                        (NaN = missing sample; the medians are precomputed per dataset as in
 *                        calculate_statistics, :123-140)
 *   nbr_off_host i32 [n_sensors+1], nbr_idx_host i32: for every sensor the sensors of its
 *                        neighbourhood in the reference's order, ITSELF INCLUDED, at most 32
 *   edge_off_host i64 [n_sensors+1], edges_host i64 [E,2]: adjacency of that neighbourhood
 *                        (distance < max_sample_distance, self-loops included) as
 *                        (receiving, sending) positions in the neighbour list
 * gnnqc_series_update overwrites the time range [t0, t0+n_t) from host slabs [n_sensors, n_t, 2].
 * gnnqc_score_cml_series: target_dev i32 [B], t_anom_dev i32 [B] -> scores_dev f32 [B]; a window
 * that the reference would skip (incomplete, or NaN in the target series) gets NaN.  n_nodes_dev
 * (optional, i32 [B]) receives the number of sensors used per window (0 = skipped). */
typedef struct gnnqc_series gnnqc_series;
int gnnqc_series_create(int device, const float* raw_host, const float* rmed_host, int64_t n_sensors,
                        int64_t n_times, const int32_t* nbr_off_host, const int32_t* nbr_idx_host,
                        const int64_t* edge_off_host, const int64_t* edges_host, gnnqc_series** out);
int gnnqc_series_update(gnnqc_series* s, int64_t t0, int64_t n_t, const float* raw_host,
                        const float* rmed_host);
/* The same, enqueued on `stream` (pinned host slabs make the copies asynchronous); the validity prefix counts are
 * rebuilt on that stream too, so a scoring call on the same stream sees the update. */
int gnnqc_series_update_async(gnnqc_series* s, int64_t t0, int64_t n_t, const float* raw_host,
                              const float* rmed_host, void* stream);
void gnnqc_series_free(gnnqc_series* s);
int gnnqc_score_cml_series(gnnqc_model* m, const gnnqc_series* series, const int32_t* target_dev,
                           const int32_t* t_anom_dev, int64_t B, int32_t timestep_before,
                           int32_t timestep_after, float* scores_dev, int32_t* n_nodes_dev, void* stream);

/* The same for SoilNet (create_tfrecords_dataset's soilnet branch, libs/preprocessing_functions.py:418-480;
 * parse_soilnet_tfrecord_fn :772-857; prepare_batch_soilnet :861-897): one window per date over ALL sensors.
 * gnnqc_soil_series_create copies to `device`:
 *   raw_host     f32 [n_sensors, n_times, 3]  moisture, temp, battv (NaN = missing sample)
 *   target_host  f32 [n_sensors, n_times]     anomaly flag (NaN = no label: the sensor is left out of the
 *                                             windows dated there, :461-464)
 *   adjacency_host u8 [n_sensors, n_sensors]  the reference's rule over the whole network,
 *       (distance <= max_sample_distance & depth difference == 0) | (distance == 0 & depth difference <=
 *       max_neighbour_depth) (:474-477); row = receiving sensor.
 * gnnqc_score_soilnet_series: t_anom_dev i32 [B] time indices -> scores_dev f32 [B, n_sensors]; a sensor the
 * reference drops from a window (no label at the date, or a NaN in one of its features inside the window), and
 * every sensor of a window that does not fit the series, gets NaN.  The window spans steps_before samples
 * before and steps_after after the date.  norm_sub3_host / norm_div3_host (both or neither): x -> (x - sub) /
 * div per feature; NULL = 'scale_range'.  valid_dev (optional) u8 [B, n_sensors]. */
typedef struct gnnqc_soil_series gnnqc_soil_series;
int gnnqc_soil_series_create(int device, const float* raw_host, const float* target_host, int64_t n_sensors,
                             int64_t n_times, const uint8_t* adjacency_host, gnnqc_soil_series** out);
void gnnqc_soil_series_free(gnnqc_soil_series* s);
int gnnqc_score_soilnet_series(gnnqc_model* m, const gnnqc_soil_series* series, const int32_t* t_anom_dev,
                               int64_t B, int32_t steps_before, int32_t steps_after,
                               const float* norm_sub3_host, const float* norm_div3_host, float* scores_dev,
                               uint8_t* valid_dev, void* stream);

/* ---- reading the reference's TFRecord window files (host only) -------------------------------- */

/* Replaces tf.data.TFRecordDataset + parse_cml_tfrecord_fn + batch + prepare_batch_cml for one file
 * written by create_tfrecords_dataset (libs/preprocessing_functions.py:388-412, schema create_example
 * :220-283, parse :567-634, batch :638-666): every tf.train.SequenceExample becomes one window of a
 * batch in the layout of gnnqc_score_cml.  normalization = the mode of parse_cml_tfrecord_fn
 * ("rolling_median" default; "standarization", "scale", "median", "rolling_median_fractional",
 * "rolling_mean", "none").  Record checksums are verified.  gnnqc_cml_records_copy fills caller
 * arrays sized from gnnqc_cml_records_info (any pointer may be NULL); _id / _dates return the target
 * sensor id and the comma-joined date strings of window i. */
typedef struct gnnqc_cml_records gnnqc_cml_records;
int gnnqc_tfrecord_read_cml(const char* path, const char* normalization, gnnqc_cml_records** out);
int gnnqc_cml_records_info(const gnnqc_cml_records* r, int64_t* n_windows, int64_t* T,
                           int64_t* n_nodes_total, int64_t* n_edges_total);
int gnnqc_cml_records_copy(const gnnqc_cml_records* r, float* features_host, float* anom_host,
                           int64_t* edges_host, int32_t* node_off_host, int64_t* edge_off_host,
                           int64_t* labels_host);
const char* gnnqc_cml_records_id(const gnnqc_cml_records* r, int64_t i);
const char* gnnqc_cml_records_dates(const gnnqc_cml_records* r, int64_t i);
void gnnqc_cml_records_free(gnnqc_cml_records* r);

/* The same for SoilNet files (create_example with ds_type == 'soilnet', :284-341; parse_soilnet_tfrecord_fn
 * :772-857; prepare_batch_soilnet :861-897): features [sum_b T*N_b, 3] = (moisture, temp, battv) in the
 * layout of gnnqc_score_soilnet, labels and sensor ids one per (window, sensor) in the order of the scores.
 * normalization: "scale_range" (default: moisture/60, temp in [-20, 40], battv in [2800, 3600]),
 * "standarization", "scale", "median" (subtraction only, as in the reference), "rolling_median",
 * "rolling_median_fractional", "rolling_mean", "none". */
typedef struct gnnqc_soilnet_records gnnqc_soilnet_records;
int gnnqc_tfrecord_read_soilnet(const char* path, const char* normalization, gnnqc_soilnet_records** out);
int gnnqc_soilnet_records_info(const gnnqc_soilnet_records* r, int64_t* n_windows, int64_t* T,
                               int64_t* n_nodes_total, int64_t* n_edges_total);
int gnnqc_soilnet_records_copy(const gnnqc_soilnet_records* r, float* features_host, int64_t* edges_host,
                               int32_t* node_off_host, int64_t* edge_off_host, int64_t* labels_host,
                               int64_t* sensor_ids_host);
const char* gnnqc_soilnet_records_dates(const gnnqc_soilnet_records* r, int64_t i);
void gnnqc_soilnet_records_free(gnnqc_soilnet_records* r);

/* ---- attribution ---------------------------------------------------------------------- */

/* Replaces IntegratedGradientsExplainer._get_gradients_single_batch for the CML GCN model
 * (xai/libs/integrated_gradients.py:1133-1234: zero baseline :898-917, interpolation :919-942,
 * per-alpha tape.gradient :955-1004, trapezoid :1006-1012, scaling by the input :1185-1191,
 * final prediction :1229).  Inputs as gnnqc_score_cml.  m_steps interpolation intervals
 * (reference default 100, xai_config_20240318.yml:12; BASELINE.json uses 50) = m_steps+1
 * forward+backward passes.
 *   attr_features_dev f32 [sum_b T*N_b, 2]  attribution of every node feature, same row order
 *                                           as features_dev (the per-sample [N_b,T,2] unwrap of
 *                                           :1017-1030 is a strided view of it)
 *   attr_anom_dev     f32 [B, T, 2]         attribution of the target series input
 *   scores_dev        f32 [B]               model output at alpha = 1
 * negative_values handling (:1193-1207) is 'keep'; abs/clip are one elementwise op for the caller. */
int gnnqc_ig_cml(gnnqc_model* m, const float* features_dev, const float* anom_dev,
                 const int64_t* edges_dev, const int32_t* node_off_dev, const int64_t* edge_off_dev,
                 int64_t B, int32_t T, int64_t n_nodes_total, int32_t m_steps,
                 float* attr_features_dev, float* attr_anom_dev, float* scores_dev, void* stream);

/* ---- evaluation on the device -------------------------------------------
 * gnnqc_select_threshold replaces select_threshold (libs/test_model.py:9-17): candidates are the
 * distinct values of round(score, 3); `score > candidate` is evaluated in float32 and the candidate with
 * the highest Matthews correlation wins, the smallest on ties.  scores_dev f32 [n] in [0, 1] (sigmoid
 * outputs; anything else, NaN included, is an error), labels_dev i64 [n] (non-zero = anomaly).
 * Outputs are HOST pointers, valid on return: threshold_out, mcc_out (optional), confusion_out
 * (optional) i64 [4] = TP, FP, TN, FN at that threshold.
 * gnnqc_roc_auc replaces sklearn.metrics.roc_auc_score as used by calculate_metrics
 * (libs/test_model.py:49; libs/metrics.py:195-211): trapezoid area under the ROC curve whose points
 * are the distinct scores; 0 when one class is absent.  n < 2^31. */
int gnnqc_select_threshold(int device, const float* scores_dev, const int64_t* labels_dev, int64_t n,
                           float* threshold_out, double* mcc_out, int64_t* confusion_out, void* stream);
int gnnqc_roc_auc(int device, const float* scores_dev, const int64_t* labels_dev, int64_t n,
                  double* auc_out, void* stream);

/* gnnqc_ig_cml without the final "(x - baseline) *" scaling: the trapezoid mean of the raw gradients, as the reference returns
 * it with integrated_gradients.scale_gradients: False (xai/libs/integrated_gradients.py:1006-1012 without :1185-1191). */
int gnnqc_ig_cml_raw(gnnqc_model* m, const float* features_dev, const float* anom_dev,
                     const int64_t* edges_dev, const int32_t* node_off_dev, const int64_t* edge_off_dev,
                     int64_t B, int32_t T, int64_t n_nodes_total, int32_t m_steps,
                     float* attr_features_dev, float* attr_anom_dev, float* scores_dev, void* stream);

/* Integrated gradients for windows named by (target sensor, time index) on a resident network (BASELINE.json config 4:
 * the IG sweep over a network that stays in HBM; reference loop prepare_data -> get_gradients ->
 * _get_gradients_single_batch, xai/libs/integrated_gradients.py:590-703,1093-1246, with the window extraction of
 * libs/preprocessing_functions.py:388-412).  Same arithmetic as gnnqc_ig_cml on the materialised windows.
 *   attr_nodes_dev f32 [B, T, 32, 2]  attribution of node n (position in the window's node list) at step t; zero padding
 *                                     beyond the window's node count, for skipped windows and for steps past the last live one
 *   attr_anom_dev  f32 [B, T, 2]      attribution of the target series input
 *   scores_dev     f32 [B]            NaN for windows the reference would skip
 *   node_sensor_dev i32 [B, 32] (optional)  global sensor id of every node slot (-1 = unused)
 *   n_nodes_dev    i32 [B] (optional) nodes per window (0 = skipped)
 * Shards by target sensor over GPUs with no collective: each rank holds its sensors' series plus their neighbours'. */
int gnnqc_ig_cml_series(gnnqc_model* m, const gnnqc_series* series, const int32_t* target_dev,
                        const int32_t* t_anom_dev, int64_t B, int32_t timestep_before,
                        int32_t timestep_after, int32_t m_steps, float* attr_nodes_dev,
                        float* attr_anom_dev, float* scores_dev, int32_t* node_sensor_dev,
                        int32_t* n_nodes_dev, void* stream);

/* ---- training ------------------------------------------------------------------------- */

/* Together these replace one train_step of model.fit as train_model sets it up
 * (libs/fit_model.py:61-112): loss 'binary_crossentropy' on the sigmoid output (Keras evaluates it
 * from the logits), class_weight {0: w0, 1: w1} as per-sample weights, mean over the batch, Adam
 * (lr; beta 0.9/0.999, epsilon 1e-7), BatchNormalization of the GeneralConv in training mode
 * (batch statistics over all node rows, moving statistics with momentum 0.99).  All four models:
 * the SoilNet GCN trains through the 'soilnet' branch of GCNClassifier.call (libs/create_model.py:224-231:
 * per-sensor neighbour mean, concat, graph_reshape) with one label per (window, sensor)
 * (libs/preprocessing_functions.py:861-897); the baselines through gnnqc_train_grads_seq.
 *
 * The step is split where data-parallel ranks must sum across GPUs - the only exchange steps:
 *   1. gnnqc_train_moments_cml: moments_dev f64[6] = {sum x0, sum x1, sum x0^2, sum x1^2,
 *      sum x0 x1, row count} of this rank's features          -> all-reduce(sum) over ranks
 *      (gnnqc_train_moments_soilnet: f64[10] = {sum x0, x1, x2, x0^2, x1^2, x2^2, x0x1, x0x2, x1x2, row count})
 *   2. gnnqc_train_grads_*: rvec_dev f32[gnnqc_train_vector_len()] = this rank's gradients in
 *      flat-parameter order, 96 slots of raw GeneralConv/BatchNorm sums (CML uses 80), and the loss in the last slot;
 *      everything is already divided by global_batch            -> all-reduce(sum) over ranks
 *   3. gnnqc_train_apply: closes the BatchNorm backward with the global sums, applies Adam to the
 *      device copy of the parameters, moves the BN statistics, refreshes the packed weights.
 * Single GPU: call 1-2-3 without the all-reduces (global_batch = B).  labels_dev: i64 [B].
 * gnnqc_train_grads_* also uses a side stream owned by the handle (weight gradients of a layer while the backward pass
 * continues below it); it forks from and joins `stream` inside the call, so for the caller everything stays ordered on
 * `stream` as usual. */
int64_t gnnqc_train_vector_len(const gnnqc_model* m);
int gnnqc_train_moments_cml(gnnqc_model* m, const float* features_dev, int64_t n_rows,
                            double* moments_dev, void* stream);
int gnnqc_train_grads_cml(gnnqc_model* m, const float* features_dev, const float* anom_dev,
                          const int64_t* edges_dev, const int32_t* node_off_dev,
                          const int64_t* edge_off_dev, const int64_t* labels_dev, int64_t B, int32_t T,
                          int64_t n_nodes_total, const double* moments_dev, float class_weight_0,
                          float class_weight_1, int64_t global_batch, float* rvec_dev, void* stream);
/* SoilNet GCNClassifier: features_dev f32 [sum N_b * T, 3] rows (window, t, sensor), edges / offsets as for
 * gnnqc_score_soilnet, labels_dev i64 [sum N_b] in (window, sensor) order; global_batch counts SEQUENCES
 * (the reference's loss is the mean over all sensor-sequences of the batch). */
int gnnqc_train_moments_soilnet(gnnqc_model* m, const float* features_dev, int64_t n_rows,
                                double* moments_dev, void* stream);
int gnnqc_train_grads_soilnet(gnnqc_model* m, const float* features_dev, const int64_t* edges_dev,
                              const int32_t* node_off_dev, const int64_t* edge_off_dev,
                              const int64_t* labels_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                              int64_t n_edges_total, const double* moments_dev, float class_weight_0,
                              float class_weight_1, int64_t global_batch, float* rvec_dev, void* stream);
/* BaselineClassifier of either dataset: x_dev f32 [n_seq, T, 2 or 3] (SoilNet: the graph_reshape'd sensor series). */
int gnnqc_train_grads_seq(gnnqc_model* m, const float* x_dev, const int64_t* labels_dev, int64_t n_seq,
                          int32_t T, float class_weight_0, float class_weight_1, int64_t global_batch,
                          float* rvec_dev, void* stream);
/* model.compile(optimizer=...) of train_model (libs/fit_model.py:71-74): 0 = Adam (default), 1 = SGD, 2 = RMSprop, all with
 * the Keras defaults and the learning rate given to gnnqc_train_apply.  Changing the rule clears the slots.
 * gnnqc_model_save writes the optimizer state the way model.save does (optimizer/_iterations, optimizer/_learning_rate,
 * optimizer/_variables/N) and gnnqc_model_load restores it, so a resumed run continues at the saved step;
 * gnnqc_train_iterations returns that step. */
int gnnqc_train_set_optimizer(gnnqc_model* m, int kind);
int64_t gnnqc_train_iterations(const gnnqc_model* m);
int gnnqc_train_apply(gnnqc_model* m, float* rvec_dev, const double* moments_dev,
                      float learning_rate, void* stream);

/* gnnqc_score_seq for callers that are not allowed to allocate.  It has no counterpart in the
 * reference.  Some users of this library cannot call cudaMalloc where they score: a
 * fixed-budget real-time loop, a process whose device memory is all owned by another
 * framework's pool, a callback that must not take allocator locks.  gnnqc_score_seq grows an
 * arena with cudaMalloc when a larger batch arrives; this variant keeps the inter-layer tiles
 * in a 32 MiB array that is part of the library image (each GPU that loads the library holds
 * those 32 MiB until the process exits), performs no cudaMalloc / cudaFree and calls no
 * allocating library routine.  Same arguments and results as gnnqc_score_seq.
 *  - Limit: the tiles of one call must fit the array - about 1,600 sequences of 181 steps or
 *    800 of 337 steps; larger calls fail with GNNQC_ERR_ARG and a message naming the limit.
 *  - The first call on a device looks up the array's device address (cudaGetSymbolAddress);
 *    make that first call before starting any stream capture.
 *  - All calls on a device share the array, which keeps the previous call's contents.  Nothing
 *    in it needs to start at zero (every tile is fully written before it is read), but calls
 *    that overlap on the GPU - other streams, handles or threads - race; serialise them.
 *  - Per-launch profiling (gnnqc_profile_enable) must be off, because it creates CUDA events. */
int gnnqc_score_seq_noalloc(gnnqc_model* m, const float* x_dev, int64_t n_seq, int32_t T,
                            float* scores_dev, void* stream);

/* The same call for windows that live in HOST memory - the shape model.predict has in the reference
 * (numpy batches from prepare_batch_cml, libs/preprocessing_functions.py:638-666, scored at
 * libs/test_model.py:29).  Arguments as gnnqc_score_cml but host pointers (pinned memory for full PCIe
 * speed); the batch is cut into chunks of `chunk_windows` (0 = 8192) that are staged to the device on a
 * copy stream while the previous chunk is scored, and scores_host [B] is complete when the call
 * returns.  Timesteps that cannot reach the score (past 162 of 181 for the shipped models) are not
 * transferred when the windows of a chunk have equal node counts.  h2d_bytes_out (optional): bytes copied
 * host-to-device by this call. */
int gnnqc_score_cml_host(gnnqc_model* m, const float* features_host, const float* anom_host,
                         const int64_t* edges_host, const int32_t* node_off_host,
                         const int64_t* edge_off_host, int64_t B, int32_t T, int64_t chunk_windows,
                         float* scores_host, int64_t* h2d_bytes_out);

/* Replaces GCNClassifier.call for ds_type 'soilnet' (libs/create_model.py:224-240), i.e.
 * model.predict on batches from prepare_batch_soilnet / wrapper_batch_mapping_soilnet
 * (libs/preprocessing_functions.py:861-897).
 *   features_dev  f32 [sum_b T*N_b, 3], rows in (window, timestep, node) order
 *   edges_dev / node_off_dev / edge_off_dev   as for gnnqc_score_cml (untiled, window-local)
 *   scores_dev    f32 [sum_b N_b]: one score per (window, sensor) in (window, node) order -
 *                 the order graph_reshape produces (libs/create_model.py:242-258) */
int gnnqc_score_soilnet(gnnqc_model* m, const float* features_dev, const int64_t* edges_dev,
                        const int32_t* node_off_dev, const int64_t* edge_off_dev, int64_t B, int32_t T,
                        int64_t n_nodes_total, int64_t n_edges_total, float* scores_dev, void* stream);

/* Bytes of device scratch the scoring entry points keep inside the handle for `n_seq`
 * sequences of length T (grown on demand with cudaMalloc, freed by gnnqc_model_free). */
int64_t gnnqc_score_scratch_bytes(const gnnqc_model* m, int64_t n_seq, int32_t T);

/* ---- measurement helpers -------------------------------------------------------------- */

/* No reference counterpart (measurement only): sustained FP32 FMA throughput of `device` in
 * TFLOP/s, mode 0 = scalar FFMA, mode 1 = packed fma.rn.f32x2.  bench.py uses it as the
 * roofline denominator of the fp32 LSTM kernels. */
int gnnqc_probe_fp32_peak(int device, int mode, double* tflops_out);

/* No reference counterpart (self-test): runs one 128x64x48 bf16 GEMM through the tcgen05 / TMEM
 * path of csrc/umma.cuh on `device` and returns its largest deviation from a host float loop. */
int gnnqc_selftest_umma(int device, float* max_abs_err);

/* No reference counterpart (measurement only).  gnnqc_launch_count: kernels this handle has
 * enqueued since it was created.  gnnqc_profile_enable(1) makes every later launch record a
 * pair of CUDA events on its stream; gnnqc_profile_read synchronises the device, sums the
 * elapsed time per kernel name into the caller's arrays (at most max_entries names of 31
 * chars) and clears the records. */
int64_t gnnqc_launch_count(const gnnqc_model* m);
int gnnqc_profile_enable(gnnqc_model* m, int on);
int gnnqc_profile_read(gnnqc_model* m, int max_entries, char (*names)[32], double* total_ms,
                       int64_t* counts, int* n_out);

#ifdef __cplusplus
}
#endif
#endif /* GNNQC_H_ */
