// TensorFlow "tensor bundle" checkpoints (variables.index + variables.data-00000-of-00001),
// read and written without TensorFlow.  Host-only code: no CUDA in this translation unit.
//
// This is the on-disk format of the SavedModel directories the reference ships
// (reference: model_cml/variables/*, loaded there by tf.keras.models.load_model,
// notebooks/pipeline.ipynb:202,436) and the format `model.save(model_path)` writes back
// (pipeline.ipynb cell 14).  The layout notes this parser follows are in SURVEY.md,
// "Appendix - TF checkpoint format".
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace gnnqc {

enum BundleDType : int { DT_FLOAT = 1, DT_INT32 = 3, DT_STRING = 7, DT_INT64 = 9 };

struct BundleEntry {
  int dtype = 0;
  std::vector<int64_t> shape;
  int64_t offset = 0;
  int64_t size = 0;
  uint32_t masked_crc = 0;
  int64_t num_elements() const {
    int64_t n = 1;
    for (int64_t d : shape) n *= d;
    return n;
  }
};

// One node of the TrackableObjectGraph stored under _CHECKPOINTABLE_OBJECT_GRAPH.
struct TrackedObject {
  std::vector<std::pair<std::string, int>> children;            // (local_name, node id)
  std::vector<std::pair<std::string, std::string>> attributes;  // (name, checkpoint_key)
  std::vector<std::string> attribute_full_names;
};

class BundleReader {
 public:
  // `prefix` is the path without the ".index" suffix, e.g. ".../variables/variables".
  // Throws std::runtime_error with a readable message on any malformed input.
  explicit BundleReader(const std::string& prefix, bool verify_crc = true);

  const std::map<std::string, BundleEntry>& entries() const { return entries_; }
  bool has(const std::string& key) const { return entries_.count(key) != 0; }
  const BundleEntry& entry(const std::string& key) const;

  std::vector<float> read_f32(const std::string& key) const;
  std::vector<int32_t> read_i32(const std::string& key) const;
  std::vector<int64_t> read_i64(const std::string& key) const;
  std::string read_scalar_string(const std::string& key) const;
  std::vector<uint8_t> read_raw(const std::string& key) const;

  // Object graph (empty if the bundle has none).
  const std::vector<TrackedObject>& objects() const { return objects_; }
  // Follow child names from the root ("time_layer/time1/cell/kernel") and return the
  // checkpoint key of that variable's VARIABLE_VALUE, or "" if the path does not exist.
  std::string key_for_path(const std::string& slash_path) const;

 private:
  void parse_object_graph();
  std::map<std::string, BundleEntry> entries_;
  std::vector<uint8_t> data_;
  std::vector<TrackedObject> objects_;
  bool verify_crc_;
};

// Writes a single-shard bundle. Tensors are laid out in key order, like TF does.
class BundleWriter {
 public:
  void add_f32(const std::string& key, const std::vector<int64_t>& shape, const float* v);
  void add_i32(const std::string& key, const std::vector<int64_t>& shape, const int32_t* v);
  void add_i64(const std::string& key, const std::vector<int64_t>& shape, const int64_t* v);
  void add_scalar_string(const std::string& key, const std::string& s);
  // Declares a trackable variable reachable from the root by `slash_path`; its value is
  // the tensor already added under `checkpoint_key`.
  void track(const std::string& slash_path, const std::string& checkpoint_key,
             const std::string& full_name);
  // Emits <prefix>.index and <prefix>.data-00000-of-00001 (object graph included when
  // track() was used).
  void finish(const std::string& prefix);

 private:
  struct Pending {
    int dtype;
    std::vector<int64_t> shape;
    std::vector<uint8_t> bytes;
  };
  std::map<std::string, Pending> tensors_;
  struct Tracked {
    std::string path, key, full_name;
  };
  std::vector<Tracked> tracked_;
};

uint32_t crc32c(const uint8_t* p, size_t n, uint32_t seed = 0);
inline uint32_t crc_mask(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xa282ead8u; }

}  // namespace gnnqc
