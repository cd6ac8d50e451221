// TFRecord reader for the reference's CML window files (host code, no TensorFlow).
//
// Replaces, for the CML GCN model, tf.data.TFRecordDataset + parse_cml_tfrecord_fn + batch +
// prepare_batch_cml (reference libs/preprocessing_functions.py:485-565, :567-634, :638-666): a file
// written by create_tfrecords_dataset (:388-412, one tf.train.SequenceExample per window, schema in
// create_example :220-283) is turned into one batch in the offsets layout of gnnqc_score_cml.
//
// File framing: per record  u64 length | u32 masked_crc32c(length) | bytes | u32 masked_crc32c(bytes).
// SequenceExample{ context = 1: Features, feature_lists = 2: FeatureLists };
// Features{ map<string, Feature> feature = 1 }, FeatureLists{ map<string, FeatureList> feature_list = 1 },
// FeatureList{ repeated Feature feature = 1 }, Feature{ bytes_list = 1 | float_list = 2 | int64_list = 3 },
// each list { repeated value = 1 } (floats / int64s packed or not).
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/gnnqc.h"
#include "capi_internal.h"
#include "tf_bundle.h"
#include "wire.h"

namespace gnnqc {
namespace {

using wire::Cursor;

[[noreturn]] void fail(const std::string& what) { throw std::runtime_error("tfrecord: " + what); }

struct Feature {
  std::vector<std::string> bytes;
  std::vector<float> floats;
  std::vector<int64_t> ints;
};

Cursor cursor_of(const std::string& s) {
  return Cursor{reinterpret_cast<const uint8_t*>(s.data()), reinterpret_cast<const uint8_t*>(s.data()) + s.size()};
}

Feature parse_feature(Cursor c) {
  Feature f;
  while (!c.done()) {
    const uint64_t tag = c.varint();
    const int field = static_cast<int>(tag >> 3), wt = static_cast<int>(tag & 7);
    if (wt != 2 || field < 1 || field > 3) { c.skip(wt); continue; }
    Cursor list = c.sub(c.varint());
    while (!list.done()) {
      const uint64_t t2 = list.varint();
      const int w2 = static_cast<int>(t2 & 7);
      if ((t2 >> 3) != 1) { list.skip(w2); continue; }
      if (field == 1 && w2 == 2) {
        Cursor b = list.sub(list.varint());
        f.bytes.emplace_back(reinterpret_cast<const char*>(b.p), b.end - b.p);
      } else if (field == 2 && w2 == 2) {          // packed floats
        Cursor b = list.sub(list.varint());
        const size_t n = static_cast<size_t>(b.end - b.p) / 4;
        const size_t at = f.floats.size();
        f.floats.resize(at + n);
        std::memcpy(f.floats.data() + at, b.p, n * 4);
      } else if (field == 2 && w2 == 5) {          // unpacked float
        const uint32_t bits = list.fixed32();
        float v;
        std::memcpy(&v, &bits, 4);
        f.floats.push_back(v);
      } else if (field == 3 && w2 == 2) {          // packed int64
        Cursor b = list.sub(list.varint());
        while (!b.done()) f.ints.push_back(static_cast<int64_t>(b.varint()));
      } else if (field == 3 && w2 == 0) {
        f.ints.push_back(static_cast<int64_t>(list.varint()));
      } else {
        list.skip(w2);
      }
    }
  }
  return f;
}

// map<string, X> entry: key = 1, value = 2
template <typename OnValue>
void parse_map(Cursor c, OnValue&& on_value) {
  while (!c.done()) {
    const uint64_t tag = c.varint();
    if ((tag >> 3) != 1 || (tag & 7) != 2) { c.skip(static_cast<int>(tag & 7)); continue; }
    Cursor entry = c.sub(c.varint());
    std::string key;
    Cursor value{nullptr, nullptr};
    while (!entry.done()) {
      const uint64_t t2 = entry.varint();
      if ((t2 & 7) != 2) { entry.skip(static_cast<int>(t2 & 7)); continue; }
      Cursor body = entry.sub(entry.varint());
      if ((t2 >> 3) == 1) key.assign(reinterpret_cast<const char*>(body.p), body.end - body.p);
      else if ((t2 >> 3) == 2) value = body;
    }
    if (value.p) on_value(key, value);
  }
}

struct SequenceExample {
  std::map<std::string, Feature> context;
  std::map<std::string, std::vector<Feature>> lists;
};

SequenceExample parse_sequence_example(const std::string& rec) {
  SequenceExample ex;
  Cursor c = cursor_of(rec);
  while (!c.done()) {
    const uint64_t tag = c.varint();
    const int field = static_cast<int>(tag >> 3), wt = static_cast<int>(tag & 7);
    if (wt != 2 || (field != 1 && field != 2)) { c.skip(wt); continue; }
    Cursor body = c.sub(c.varint());
    if (field == 1) {
      parse_map(body, [&](const std::string& key, Cursor v) { ex.context[key] = parse_feature(v); });
    } else {
      parse_map(body, [&](const std::string& key, Cursor v) {
        std::vector<Feature>& out = ex.lists[key];
        while (!v.done()) {
          const uint64_t t2 = v.varint();
          if ((t2 >> 3) == 1 && (t2 & 7) == 2) out.push_back(parse_feature(v.sub(v.varint())));
          else v.skip(static_cast<int>(t2 & 7));
        }
      });
    }
  }
  return ex;
}

std::vector<std::string> read_records(const std::string& path, bool verify) {
  std::ifstream f(path, std::ios::binary);
  if (!f) fail("cannot open " + path);
  std::vector<std::string> out;
  while (true) {
    uint8_t head[12];
    f.read(reinterpret_cast<char*>(head), 12);
    if (f.gcount() == 0) break;
    if (f.gcount() != 12) fail("truncated record header in " + path);
    uint64_t len;
    uint32_t len_crc;
    std::memcpy(&len, head, 8);
    std::memcpy(&len_crc, head + 8, 4);
    if (verify && crc_mask(crc32c(head, 8)) != len_crc) fail("record length checksum mismatch in " + path);
    if (len > (uint64_t(1) << 31)) fail("implausible record length in " + path);
    std::string data(static_cast<size_t>(len), '\0');
    f.read(&data[0], static_cast<std::streamsize>(len));
    uint32_t data_crc;
    f.read(reinterpret_cast<char*>(&data_crc), 4);
    if (!f) fail("truncated record in " + path);
    if (verify && crc_mask(crc32c(reinterpret_cast<const uint8_t*>(data.data()), data.size())) != data_crc)
      fail("record data checksum mismatch in " + path);
    out.push_back(std::move(data));
  }
  return out;
}


// One parsed file = one batch in the offsets layout.
struct CmlBatch {
  int64_t T = 0;
  std::vector<float> features, anom;
  std::vector<int64_t> edges, edge_off, labels;
  std::vector<int32_t> node_off;
  std::vector<std::string> ids, dates;   // target sensor id and the window's date strings joined by ','
};

const Feature& need(const std::map<std::string, Feature>& m, const std::string& key) {
  auto it = m.find(key);
  if (it == m.end()) fail("SequenceExample has no context feature '" + key + "'");
  return it->second;
}

// parse_cml_tfrecord_fn's normalisation modes (reference :606-625): out = (x - sub[n]) / div[n].
void normaliser(const SequenceExample& ex, const std::string& mode, int ch, size_t n_nodes, std::vector<float>* sub, std::vector<float>* div) {
  const std::string p = ch == 0 ? "TRSL1_" : "TRSL2_";
  sub->assign(n_nodes, 0.f);
  div->assign(n_nodes, 1.f);
  auto load = [&](const std::string& key, std::vector<float>* dst) {
    const Feature& f = need(ex.context, p + key);
    if (f.floats.size() != n_nodes) fail("context feature '" + p + key + "' has the wrong length");
    *dst = f.floats;
  };
  if (mode == "rolling_median") load("rolling_median", sub);
  else if (mode == "standarization") { load("mean", sub); load("std", div); }
  else if (mode == "median") { load("median", sub); load("median", div); }
  else if (mode == "rolling_median_fractional") { load("rolling_median", sub); load("rolling_median", div); }
  else if (mode == "rolling_mean") { load("rolling_mean", sub); load("rolling_std", div); }
  else if (mode == "scale") {
    std::vector<float> mx;
    load("min", sub);
    load("max", &mx);
    for (size_t i = 0; i < n_nodes; ++i) (*div)[i] = mx[i] - (*sub)[i];
  } else if (mode != "none" && !mode.empty()) fail("unknown normalisation '" + mode + "'");
}

void append_example(const SequenceExample& ex, const std::string& mode, CmlBatch* b) {
  auto l1 = ex.lists.find("TRSL1"), l2 = ex.lists.find("TRSL2");
  if (l1 == ex.lists.end() || l2 == ex.lists.end()) fail("SequenceExample has no TRSL1/TRSL2 feature lists");
  const size_t T = l1->second.size();
  if (T == 0 || l2->second.size() != T) fail("TRSL1/TRSL2 have different lengths");
  if (b->T == 0) b->T = static_cast<int64_t>(T);
  if (static_cast<size_t>(b->T) != T) fail("windows of different lengths in one file");
  const size_t N = l1->second[0].floats.size();
  if (N == 0) fail("window without sensors");
  const Feature& ids = need(ex.context, "CML_ids");
  const Feature& target = need(ex.context, "anomaly_ID");
  if (ids.bytes.size() != N || target.bytes.size() != 1) fail("CML_ids / anomaly_ID malformed");
  size_t tgt = N;
  for (size_t i = 0; i < N; ++i)
    if (ids.bytes[i] == target.bytes[0]) { tgt = i; break; }
  if (tgt == N) fail("anomaly_ID is not among CML_ids");
  std::vector<float> sub[2], div[2];
  normaliser(ex, mode, 0, N, &sub[0], &div[0]);
  normaliser(ex, mode, 1, N, &sub[1], &div[1]);
  const size_t f0 = b->features.size(), a0 = b->anom.size();
  b->features.resize(f0 + T * N * 2);
  b->anom.resize(a0 + T * 2);
  for (size_t t = 0; t < T; ++t) {
    const std::vector<float>& r1 = l1->second[t].floats;
    const std::vector<float>& r2 = l2->second[t].floats;
    if (r1.size() != N || r2.size() != N) fail("ragged timestep inside a window");
    for (size_t n = 0; n < N; ++n) {
      const float v1 = (r1[n] - sub[0][n]) / div[0][n], v2 = (r2[n] - sub[1][n]) / div[1][n];
      b->features[f0 + (t * N + n) * 2] = v1;
      b->features[f0 + (t * N + n) * 2 + 1] = v2;
      if (n == tgt) { b->anom[a0 + t * 2] = v1; b->anom[a0 + t * 2 + 1] = v2; }
    }
  }
  auto nodes = ex.lists.find("nodes"), nbrs = ex.lists.find("neighbours");
  if (nodes == ex.lists.end() || nbrs == ex.lists.end() || nodes->second.size() != nbrs->second.size()) fail("nodes / neighbours malformed");
  for (size_t e = 0; e < nodes->second.size(); ++e) {
    if (nodes->second[e].ints.size() != 1 || nbrs->second[e].ints.size() != 1) fail("edge feature without exactly one index");
    b->edges.push_back(nodes->second[e].ints[0]);
    b->edges.push_back(nbrs->second[e].ints[0]);
  }
  b->node_off.push_back(b->node_off.back() + static_cast<int32_t>(N));
  b->edge_off.push_back(b->edge_off.back() + static_cast<int64_t>(nodes->second.size()));
  const Feature& flag = need(ex.context, "anomaly_flag");
  b->labels.push_back(flag.ints.empty() ? 0 : flag.ints[0]);
  b->ids.push_back(target.bytes[0]);
  std::string dates;
  auto d = ex.context.find("dates");
  if (d != ex.context.end())
    for (size_t i = 0; i < d->second.bytes.size(); ++i) dates += (i ? "," : "") + d->second.bytes[i];
  b->dates.push_back(dates);
}

// ---- SoilNet windows (create_example, ds_type == 'soilnet'; parse_soilnet_tfrecord_fn) ---------------
// One parsed file = one batch in the offsets layout of gnnqc_score_soilnet: features [sum T*N_b, 3]
// (moisture, temp, battv), one label and one sensor id per (window, sensor).
struct SoilBatch {
  int64_t T = 0;
  std::vector<float> features;
  std::vector<int64_t> edges, edge_off, labels, sensor_ids;
  std::vector<int32_t> node_off;
  std::vector<std::string> dates;
};

// parse_soilnet_tfrecord_fn's normalisation modes (reference :821-852): out = (x - sub[n]) / div[n].
void soil_normaliser(const SequenceExample& ex, const std::string& mode, int ch, size_t n_nodes, std::vector<float>* sub,
                     std::vector<float>* div) {
  static const char* const kNames[3] = {"moisture_", "temp_", "battv_"};
  const std::string p = kNames[ch];
  sub->assign(n_nodes, 0.f);
  div->assign(n_nodes, 1.f);
  auto load = [&](const std::string& key, std::vector<float>* dst) {
    const Feature& f = need(ex.context, p + key);
    if (f.floats.size() != n_nodes) fail("context feature '" + p + key + "' has the wrong length");
    *dst = f.floats;
  };
  if (mode == "scale_range") {   // fixed physical ranges: moisture / 60, temp in [-20, 40], battv in [2800, 3600]
    static const float kSub[3] = {0.f, -20.f, 2800.f}, kDiv[3] = {60.f, 60.f, 800.f};
    sub->assign(n_nodes, kSub[ch]);
    div->assign(n_nodes, kDiv[ch]);
  } else if (mode == "standarization") { load("mean", sub); load("std", div); }
  else if (mode == "median") load("median", sub);                 // the reference's division is commented out
  else if (mode == "rolling_median") load("rolling_median", sub);
  else if (mode == "rolling_median_fractional") { load("rolling_median", sub); load("rolling_median", div); }
  else if (mode == "rolling_mean") { load("rolling_mean", sub); load("rolling_std", div); }
  else if (mode == "scale") {
    std::vector<float> mx;
    load("min", sub);
    load("max", &mx);
    for (size_t i = 0; i < n_nodes; ++i) (*div)[i] = mx[i] - (*sub)[i];
  } else if (mode != "none" && !mode.empty()) fail("unknown normalisation '" + mode + "'");
}

const std::vector<Feature>& need_list(const SequenceExample& ex, const std::string& key) {
  auto it = ex.lists.find(key);
  if (it == ex.lists.end()) fail("SequenceExample has no feature list '" + key + "'");
  return it->second;
}

void append_soil_example(const SequenceExample& ex, const std::string& mode, SoilBatch* b) {
  static const char* const kLists[3] = {"moisture", "temp", "battv"};
  const std::vector<Feature>* ch[3];
  for (int c = 0; c < 3; ++c) ch[c] = &need_list(ex, kLists[c]);
  const size_t T = ch[0]->size();
  if (T == 0 || ch[1]->size() != T || ch[2]->size() != T) fail("moisture / temp / battv have different lengths");
  if (b->T == 0) b->T = static_cast<int64_t>(T);
  if (static_cast<size_t>(b->T) != T) fail("windows of different lengths in one file");
  const size_t N = (*ch[0])[0].floats.size();
  if (N == 0) fail("window without sensors");
  std::vector<float> sub[3], div[3];
  for (int c = 0; c < 3; ++c) soil_normaliser(ex, mode, c, N, &sub[c], &div[c]);
  const size_t f0 = b->features.size();
  b->features.resize(f0 + T * N * 3);
  for (size_t t = 0; t < T; ++t)
    for (int c = 0; c < 3; ++c) {
      const std::vector<float>& row = (*ch[c])[t].floats;
      if (row.size() != N) fail("ragged timestep inside a window");
      for (size_t n = 0; n < N; ++n) b->features[f0 + (t * N + n) * 3 + c] = (row[n] - sub[c][n]) / div[c][n];
    }
  const std::vector<Feature>& nodes = need_list(ex, "nodes");
  const std::vector<Feature>& nbrs = need_list(ex, "neighbours");
  if (nodes.size() != nbrs.size()) fail("nodes / neighbours malformed");
  for (size_t e = 0; e < nodes.size(); ++e) {
    if (nodes[e].ints.size() != 1 || nbrs[e].ints.size() != 1) fail("edge feature without exactly one index");
    b->edges.push_back(nodes[e].ints[0]);
    b->edges.push_back(nbrs[e].ints[0]);
  }
  const std::vector<Feature>& flags = need_list(ex, "anomaly_flag");
  const std::vector<Feature>& sids = need_list(ex, "sensor_ids");
  if (flags.size() != N || sids.size() != N) fail("anomaly_flag / sensor_ids do not have one entry per sensor");
  for (size_t n = 0; n < N; ++n) {
    if (flags[n].ints.size() != 1 || sids[n].ints.size() != 1) fail("anomaly_flag / sensor_ids feature without exactly one value");
    b->labels.push_back(flags[n].ints[0]);
    b->sensor_ids.push_back(sids[n].ints[0]);
  }
  b->node_off.push_back(b->node_off.back() + static_cast<int32_t>(N));
  b->edge_off.push_back(b->edge_off.back() + static_cast<int64_t>(nodes.size()));
  std::string dates;
  auto d = ex.context.find("dates");
  if (d != ex.context.end())
    for (size_t i = 0; i < d->second.bytes.size(); ++i) dates += (i ? "," : "") + d->second.bytes[i];
  b->dates.push_back(dates);
}

}  // namespace
}  // namespace gnnqc

struct gnnqc_soilnet_records {
  gnnqc::SoilBatch batch;
};

struct gnnqc_cml_records {
  gnnqc::CmlBatch batch;
};

extern "C" {

int gnnqc_tfrecord_read_cml(const char* path, const char* normalization, gnnqc_cml_records** out) {
  if (!path || !out) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_tfrecord_read_cml: null argument");
  *out = nullptr;
  gnnqc_cml_records* r = new gnnqc_cml_records();
  try {
    r->batch.node_off.push_back(0);
    r->batch.edge_off.push_back(0);
    const std::string mode = normalization ? normalization : "rolling_median";
    for (const std::string& rec : gnnqc::read_records(path, true)) gnnqc::append_example(gnnqc::parse_sequence_example(rec), mode, &r->batch);
  } catch (const std::exception& e) {
    delete r;
    return gnnqc_internal_fail(GNNQC_ERR_CKPT, e.what());
  }
  *out = r;
  return GNNQC_OK;
}

int gnnqc_cml_records_info(const gnnqc_cml_records* r, int64_t* n_windows, int64_t* T, int64_t* n_nodes_total, int64_t* n_edges_total) {
  if (!r) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_cml_records_info: null handle");
  if (n_windows) *n_windows = static_cast<int64_t>(r->batch.labels.size());
  if (T) *T = r->batch.T;
  if (n_nodes_total) *n_nodes_total = r->batch.node_off.back();
  if (n_edges_total) *n_edges_total = r->batch.edge_off.back();
  return GNNQC_OK;
}

int gnnqc_cml_records_copy(const gnnqc_cml_records* r, float* features_host, float* anom_host, int64_t* edges_host, int32_t* node_off_host,
                           int64_t* edge_off_host, int64_t* labels_host) {
  if (!r) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_cml_records_copy: null handle");
  const gnnqc::CmlBatch& b = r->batch;
  if (features_host) std::memcpy(features_host, b.features.data(), b.features.size() * sizeof(float));
  if (anom_host) std::memcpy(anom_host, b.anom.data(), b.anom.size() * sizeof(float));
  if (edges_host) std::memcpy(edges_host, b.edges.data(), b.edges.size() * sizeof(int64_t));
  if (node_off_host) std::memcpy(node_off_host, b.node_off.data(), b.node_off.size() * sizeof(int32_t));
  if (edge_off_host) std::memcpy(edge_off_host, b.edge_off.data(), b.edge_off.size() * sizeof(int64_t));
  if (labels_host) std::memcpy(labels_host, b.labels.data(), b.labels.size() * sizeof(int64_t));
  return GNNQC_OK;
}

const char* gnnqc_cml_records_id(const gnnqc_cml_records* r, int64_t i) {
  return (r && i >= 0 && i < static_cast<int64_t>(r->batch.ids.size())) ? r->batch.ids[static_cast<size_t>(i)].c_str() : "";
}

const char* gnnqc_cml_records_dates(const gnnqc_cml_records* r, int64_t i) {
  return (r && i >= 0 && i < static_cast<int64_t>(r->batch.dates.size())) ? r->batch.dates[static_cast<size_t>(i)].c_str() : "";
}

void gnnqc_cml_records_free(gnnqc_cml_records* r) { delete r; }

int gnnqc_tfrecord_read_soilnet(const char* path, const char* normalization, gnnqc_soilnet_records** out) {
  if (!path || !out) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_tfrecord_read_soilnet: null argument");
  *out = nullptr;
  gnnqc_soilnet_records* r = new gnnqc_soilnet_records();
  try {
    r->batch.node_off.push_back(0);
    r->batch.edge_off.push_back(0);
    const std::string mode = normalization ? normalization : "scale_range";
    for (const std::string& rec : gnnqc::read_records(path, true))
      gnnqc::append_soil_example(gnnqc::parse_sequence_example(rec), mode, &r->batch);
  } catch (const std::exception& e) {
    delete r;
    return gnnqc_internal_fail(GNNQC_ERR_CKPT, e.what());
  }
  *out = r;
  return GNNQC_OK;
}

int gnnqc_soilnet_records_info(const gnnqc_soilnet_records* r, int64_t* n_windows, int64_t* T, int64_t* n_nodes_total, int64_t* n_edges_total) {
  if (!r) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_soilnet_records_info: null handle");
  if (n_windows) *n_windows = static_cast<int64_t>(r->batch.dates.size());
  if (T) *T = r->batch.T;
  if (n_nodes_total) *n_nodes_total = r->batch.node_off.back();
  if (n_edges_total) *n_edges_total = r->batch.edge_off.back();
  return GNNQC_OK;
}

int gnnqc_soilnet_records_copy(const gnnqc_soilnet_records* r, float* features_host, int64_t* edges_host, int32_t* node_off_host,
                               int64_t* edge_off_host, int64_t* labels_host, int64_t* sensor_ids_host) {
  if (!r) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_soilnet_records_copy: null handle");
  const gnnqc::SoilBatch& b = r->batch;
  if (features_host) std::memcpy(features_host, b.features.data(), b.features.size() * sizeof(float));
  if (edges_host) std::memcpy(edges_host, b.edges.data(), b.edges.size() * sizeof(int64_t));
  if (node_off_host) std::memcpy(node_off_host, b.node_off.data(), b.node_off.size() * sizeof(int32_t));
  if (edge_off_host) std::memcpy(edge_off_host, b.edge_off.data(), b.edge_off.size() * sizeof(int64_t));
  if (labels_host) std::memcpy(labels_host, b.labels.data(), b.labels.size() * sizeof(int64_t));
  if (sensor_ids_host) std::memcpy(sensor_ids_host, b.sensor_ids.data(), b.sensor_ids.size() * sizeof(int64_t));
  return GNNQC_OK;
}

const char* gnnqc_soilnet_records_dates(const gnnqc_soilnet_records* r, int64_t i) {
  return (r && i >= 0 && i < static_cast<int64_t>(r->batch.dates.size())) ? r->batch.dates[static_cast<size_t>(i)].c_str() : "";
}

void gnnqc_soilnet_records_free(gnnqc_soilnet_records* r) { delete r; }

}  // extern "C"
