// See tf_bundle.h.  Format facts (verified against the four shipped checkpoints):
//  * variables.index is a LevelDB-style table: data blocks, metaindex block, index block,
//    48-byte footer ending in the magic 0xdb4775248b80fb57.  Every block is followed by a
//    1-byte compression type (0 = raw, the only type accepted here) and a masked crc32c of
//    block+type.
//  * key ""  -> BundleHeaderProto; every other key -> BundleEntryProto.
//  * tensors live in the .data file at [offset, offset+size); entry.crc32c is the masked
//    crc32c of those bytes.  A string tensor is: varint length per element, masked crc32c
//    of the lengths (as uint32 each), then the bytes.
#include "tf_bundle.h"
#include "wire.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>

namespace gnnqc {
namespace {

constexpr uint64_t kTableMagic = 0xdb4775248b80fb57ull;
const char* const kObjectGraphKey = "_CHECKPOINTABLE_OBJECT_GRAPH";

[[noreturn]] void fail(const std::string& what) { throw std::runtime_error("tf_bundle: " + what); }
using wire::Cursor;

std::vector<uint8_t> slurp(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) fail("cannot open " + path);
  f.seekg(0, std::ios::end);
  std::streamoff n = f.tellg();
  f.seekg(0);
  std::vector<uint8_t> buf(static_cast<size_t>(n));
  if (n > 0 && !f.read(reinterpret_cast<char*>(buf.data()), n)) fail("short read on " + path);
  return buf;
}

struct BlockHandle {
  uint64_t offset, size;
};

BlockHandle read_handle(Cursor& c) {
  BlockHandle h;
  h.offset = c.varint();
  h.size = c.varint();
  return h;
}

// Decodes one table block into (key, value) pairs.
void read_block(const std::vector<uint8_t>& file, BlockHandle h, bool verify,
                std::vector<std::pair<std::string, std::string>>* out) {
  if (h.offset + h.size + 5 > file.size()) fail("block handle outside the index file");
  const uint8_t* base = file.data() + h.offset;
  if (base[h.size] != 0) fail("compressed index block (type " + std::to_string(base[h.size]) + ") is not supported");
  if (verify) {
    uint32_t want;
    std::memcpy(&want, base + h.size + 1, 4);
    if (crc_mask(crc32c(base, h.size + 1)) != want) fail("index block checksum mismatch");
  }
  if (h.size < 4) fail("index block too small");
  uint32_t n_restarts;
  std::memcpy(&n_restarts, base + h.size - 4, 4);
  uint64_t tail = 4ull + 4ull * n_restarts;
  if (tail > h.size) fail("bad restart array");
  Cursor c{base, base + (h.size - tail)};
  std::string key;
  while (!c.done()) {
    uint64_t shared = c.varint(), fresh = c.varint(), vlen = c.varint();
    if (shared > key.size()) fail("bad shared-prefix length");
    Cursor k = c.sub(fresh);
    key.resize(shared);
    key.append(reinterpret_cast<const char*>(k.p), fresh);
    Cursor v = c.sub(vlen);
    out->emplace_back(key, std::string(reinterpret_cast<const char*>(v.p), vlen));
  }
}

BundleEntry decode_entry(const std::string& bytes) {
  BundleEntry e;
  Cursor c{reinterpret_cast<const uint8_t*>(bytes.data()),
           reinterpret_cast<const uint8_t*>(bytes.data()) + bytes.size()};
  while (!c.done()) {
    uint64_t tag = c.varint();
    int field = static_cast<int>(tag >> 3), wire = static_cast<int>(tag & 7);
    if (field == 1 && wire == 0) e.dtype = static_cast<int>(c.varint());
    else if (field == 2 && wire == 2) {          // TensorShapeProto
      Cursor s = c.sub(c.varint());
      while (!s.done()) {
        uint64_t t2 = s.varint();
        if ((t2 >> 3) == 2 && (t2 & 7) == 2) {   // repeated Dim
          Cursor d = s.sub(s.varint());
          int64_t sz = 0;
          while (!d.done()) {
            uint64_t t3 = d.varint();
            if ((t3 >> 3) == 1 && (t3 & 7) == 0) sz = static_cast<int64_t>(d.varint());
            else d.skip(static_cast<int>(t3 & 7));
          }
          e.shape.push_back(sz);
        } else s.skip(static_cast<int>(t2 & 7));
      }
    } else if (field == 3 && wire == 0) {
      if (c.varint() != 0) fail("multi-shard bundles are not supported");
    } else if (field == 4 && wire == 0) e.offset = static_cast<int64_t>(c.varint());
    else if (field == 5 && wire == 0) e.size = static_cast<int64_t>(c.varint());
    else if (field == 6 && wire == 5) e.masked_crc = c.fixed32();
    else c.skip(wire);
  }
  return e;
}

// ---- tiny protobuf emitter -------------------------------------------------------------
void put_varint(std::string* s, uint64_t v) {
  while (v >= 0x80) {
    s->push_back(static_cast<char>((v & 0x7f) | 0x80));
    v >>= 7;
  }
  s->push_back(static_cast<char>(v));
}
void put_tag(std::string* s, int field, int wire) { put_varint(s, (static_cast<uint64_t>(field) << 3) | wire); }
void put_bytes(std::string* s, int field, const std::string& b) {
  put_tag(s, field, 2);
  put_varint(s, b.size());
  s->append(b);
}
void put_u32(std::string* s, uint32_t v) { s->append(reinterpret_cast<const char*>(&v), 4); }

std::vector<std::string> split_path(const std::string& p) {
  std::vector<std::string> parts;
  size_t i = 0;
  while (i <= p.size()) {
    size_t j = p.find('/', i);
    if (j == std::string::npos) j = p.size();
    if (j > i) parts.push_back(p.substr(i, j - i));
    i = j + 1;
  }
  return parts;
}

}  // namespace

// ---- crc32c (Castagnoli), slicing-by-1 ----------------------------------------------------
uint32_t crc32c(const uint8_t* p, size_t n, uint32_t seed) {
  static uint32_t table[256];
  static bool ready = false;
  if (!ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1) ? 0x82F63B78u : 0u);
      table[i] = c;
    }
    ready = true;
  }
  uint32_t c = ~seed;
  for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xff] ^ (c >> 8);
  return ~c;
}

// ---- reader -------------------------------------------------------------------------------
BundleReader::BundleReader(const std::string& prefix, bool verify_crc) : verify_crc_(verify_crc) {
  std::vector<uint8_t> index = slurp(prefix + ".index");
  if (index.size() < 48) fail("index file too small");
  const uint8_t* foot = index.data() + index.size() - 48;
  uint64_t magic;
  std::memcpy(&magic, foot + 40, 8);
  if (magic != kTableMagic) fail("bad table magic in " + prefix + ".index");
  Cursor fc{foot, foot + 40};
  read_handle(fc);  // metaindex: unused by tensor bundles
  BlockHandle index_block = read_handle(fc);

  std::vector<std::pair<std::string, std::string>> idx, kv;
  read_block(index, index_block, verify_crc_, &idx);
  for (auto& ent : idx) {
    Cursor hc{reinterpret_cast<const uint8_t*>(ent.second.data()),
              reinterpret_cast<const uint8_t*>(ent.second.data()) + ent.second.size()};
    read_block(index, read_handle(hc), verify_crc_, &kv);
  }
  bool saw_header = false;
  for (auto& p : kv) {
    if (p.first.empty()) {
      // BundleHeaderProto{num_shards=1, endianness=2, version=3}
      Cursor hc{reinterpret_cast<const uint8_t*>(p.second.data()),
                reinterpret_cast<const uint8_t*>(p.second.data()) + p.second.size()};
      while (!hc.done()) {
        uint64_t tag = hc.varint();
        if ((tag >> 3) == 1 && (tag & 7) == 0) {
          if (hc.varint() != 1) fail("only single-shard bundles are supported");
        } else if ((tag >> 3) == 2 && (tag & 7) == 0) {
          if (hc.varint() != 0) fail("big-endian bundles are not supported");
        } else hc.skip(static_cast<int>(tag & 7));
      }
      saw_header = true;
    } else {
      entries_[p.first] = decode_entry(p.second);
    }
  }
  if (!saw_header) fail("bundle header entry missing");
  data_ = slurp(prefix + ".data-00000-of-00001");
  for (auto& kvp : entries_) {
    const BundleEntry& e = kvp.second;
    if (e.offset < 0 || e.size < 0 || static_cast<uint64_t>(e.offset + e.size) > data_.size())
      fail("tensor '" + kvp.first + "' lies outside the data file");
    if (verify_crc_ && e.dtype != DT_STRING &&
        crc_mask(crc32c(data_.data() + e.offset, static_cast<size_t>(e.size))) != e.masked_crc)
      fail("tensor '" + kvp.first + "' checksum mismatch");
  }
  parse_object_graph();
}

const BundleEntry& BundleReader::entry(const std::string& key) const {
  auto it = entries_.find(key);
  if (it == entries_.end()) fail("no tensor named '" + key + "'");
  return it->second;
}

std::vector<uint8_t> BundleReader::read_raw(const std::string& key) const {
  const BundleEntry& e = entry(key);
  return std::vector<uint8_t>(data_.begin() + e.offset, data_.begin() + e.offset + e.size);
}

template <typename T>
static std::vector<T> read_typed(const BundleReader& r, const std::string& key, int dtype) {
  const BundleEntry& e = r.entry(key);
  if (e.dtype != dtype) fail("tensor '" + key + "' has dtype " + std::to_string(e.dtype));
  if (e.size != e.num_elements() * static_cast<int64_t>(sizeof(T))) fail("tensor '" + key + "' size/shape mismatch");
  std::vector<uint8_t> raw = r.read_raw(key);
  std::vector<T> out(static_cast<size_t>(e.num_elements()));
  if (!out.empty()) std::memcpy(out.data(), raw.data(), raw.size());
  return out;
}
std::vector<float> BundleReader::read_f32(const std::string& key) const { return read_typed<float>(*this, key, DT_FLOAT); }
std::vector<int32_t> BundleReader::read_i32(const std::string& key) const { return read_typed<int32_t>(*this, key, DT_INT32); }
std::vector<int64_t> BundleReader::read_i64(const std::string& key) const { return read_typed<int64_t>(*this, key, DT_INT64); }

std::string BundleReader::read_scalar_string(const std::string& key) const {
  const BundleEntry& e = entry(key);
  if (e.dtype != DT_STRING || e.num_elements() != 1) fail("tensor '" + key + "' is not a scalar string");
  Cursor c{data_.data() + e.offset, data_.data() + e.offset + e.size};
  uint64_t len = c.varint();
  uint32_t len_crc = c.fixed32();
  if (verify_crc_) {
    uint32_t l32 = static_cast<uint32_t>(len);
    if (crc_mask(crc32c(reinterpret_cast<const uint8_t*>(&l32), 4)) != len_crc) fail("string length checksum mismatch in '" + key + "'");
  }
  Cursor body = c.sub(len);
  return std::string(reinterpret_cast<const char*>(body.p), len);
}

void BundleReader::parse_object_graph() {
  if (!has(kObjectGraphKey)) return;
  std::string blob = read_scalar_string(kObjectGraphKey);
  Cursor g{reinterpret_cast<const uint8_t*>(blob.data()), reinterpret_cast<const uint8_t*>(blob.data()) + blob.size()};
  while (!g.done()) {
    uint64_t tag = g.varint();
    if ((tag >> 3) != 1 || (tag & 7) != 2) {  // only `repeated TrackableObject nodes = 1`
      g.skip(static_cast<int>(tag & 7));
      continue;
    }
    Cursor n = g.sub(g.varint());
    TrackedObject obj;
    while (!n.done()) {
      uint64_t t = n.varint();
      int field = static_cast<int>(t >> 3), wire = static_cast<int>(t & 7);
      if (field == 1 && wire == 2) {  // ObjectReference{node_id=1, local_name=2}
        Cursor r = n.sub(n.varint());
        int id = 0;
        std::string name;
        while (!r.done()) {
          uint64_t t2 = r.varint();
          if ((t2 >> 3) == 1 && (t2 & 7) == 0) id = static_cast<int>(r.varint());
          else if ((t2 >> 3) == 2 && (t2 & 7) == 2) {
            Cursor s = r.sub(r.varint());
            name.assign(reinterpret_cast<const char*>(s.p), s.end - s.p);
          } else r.skip(static_cast<int>(t2 & 7));
        }
        obj.children.emplace_back(name, id);
      } else if (field == 2 && wire == 2) {  // SerializedTensor{name=1, full_name=2, checkpoint_key=3}
        Cursor a = n.sub(n.varint());
        std::string name, full, key;
        while (!a.done()) {
          uint64_t t2 = a.varint();
          int f2 = static_cast<int>(t2 >> 3);
          if ((t2 & 7) == 2 && f2 >= 1 && f2 <= 3) {
            Cursor s = a.sub(a.varint());
            std::string v(reinterpret_cast<const char*>(s.p), s.end - s.p);
            (f2 == 1 ? name : f2 == 2 ? full : key) = v;
          } else a.skip(static_cast<int>(t2 & 7));
        }
        obj.attributes.emplace_back(name, key);
        obj.attribute_full_names.push_back(full);
      } else n.skip(wire);
    }
    objects_.push_back(std::move(obj));
  }
}

std::string BundleReader::key_for_path(const std::string& slash_path) const {
  if (objects_.empty()) return "";
  int node = 0;
  for (const std::string& part : split_path(slash_path)) {
    int next = -1;
    for (auto& ch : objects_[node].children)
      if (ch.first == part) { next = ch.second; break; }
    if (next < 0 || next >= static_cast<int>(objects_.size())) return "";
    node = next;
  }
  for (auto& at : objects_[node].attributes)
    if (at.first == "VARIABLE_VALUE") return at.second;
  return "";
}

// ---- writer -------------------------------------------------------------------------------
void BundleWriter::add_f32(const std::string& key, const std::vector<int64_t>& shape, const float* v) {
  Pending p{DT_FLOAT, shape, {}};
  int64_t n = 1;
  for (int64_t d : shape) n *= d;
  p.bytes.assign(reinterpret_cast<const uint8_t*>(v), reinterpret_cast<const uint8_t*>(v) + n * 4);
  tensors_[key] = std::move(p);
}
void BundleWriter::add_i32(const std::string& key, const std::vector<int64_t>& shape, const int32_t* v) {
  Pending p{DT_INT32, shape, {}};
  int64_t n = 1;
  for (int64_t d : shape) n *= d;
  p.bytes.assign(reinterpret_cast<const uint8_t*>(v), reinterpret_cast<const uint8_t*>(v) + n * 4);
  tensors_[key] = std::move(p);
}
void BundleWriter::add_i64(const std::string& key, const std::vector<int64_t>& shape, const int64_t* v) {
  Pending p{DT_INT64, shape, {}};
  int64_t n = 1;
  for (int64_t d : shape) n *= d;
  p.bytes.assign(reinterpret_cast<const uint8_t*>(v), reinterpret_cast<const uint8_t*>(v) + n * 8);
  tensors_[key] = std::move(p);
}
void BundleWriter::add_scalar_string(const std::string& key, const std::string& s) {
  Pending p{DT_STRING, {}, {}};
  std::string enc;
  put_varint(&enc, s.size());
  uint32_t l32 = static_cast<uint32_t>(s.size());
  put_u32(&enc, crc_mask(crc32c(reinterpret_cast<const uint8_t*>(&l32), 4)));
  enc.append(s);
  p.bytes.assign(enc.begin(), enc.end());
  tensors_[key] = std::move(p);
}
void BundleWriter::track(const std::string& slash_path, const std::string& checkpoint_key, const std::string& full_name) {
  tracked_.push_back({slash_path, checkpoint_key, full_name});
}

void BundleWriter::finish(const std::string& prefix) {
  if (!tracked_.empty()) {
    // Build the object tree from the tracked paths, then serialise it as a
    // TrackableObjectGraph (nodes in creation order, root = node 0).
    std::vector<TrackedObject> nodes(1);
    for (const Tracked& t : tracked_) {
      int node = 0;
      for (const std::string& part : split_path(t.path)) {
        int next = -1;
        for (auto& ch : nodes[node].children)
          if (ch.first == part) next = ch.second;
        if (next < 0) {
          next = static_cast<int>(nodes.size());
          nodes[node].children.emplace_back(part, next);
          nodes.emplace_back();
        }
        node = next;
      }
      nodes[node].attributes.emplace_back("VARIABLE_VALUE", t.key);
      nodes[node].attribute_full_names.push_back(t.full_name);
    }
    std::string graph;
    for (const TrackedObject& o : nodes) {
      std::string n;
      for (auto& ch : o.children) {
        std::string r;
        put_tag(&r, 1, 0);
        put_varint(&r, static_cast<uint64_t>(ch.second));
        put_bytes(&r, 2, ch.first);
        put_bytes(&n, 1, r);
      }
      for (size_t i = 0; i < o.attributes.size(); ++i) {
        std::string a;
        put_bytes(&a, 1, o.attributes[i].first);
        put_bytes(&a, 2, o.attribute_full_names[i]);
        put_bytes(&a, 3, o.attributes[i].second);
        put_bytes(&n, 2, a);
      }
      put_bytes(&graph, 1, n);
    }
    add_scalar_string(kObjectGraphKey, graph);
  }

  // Data file + entry protos, in key order.
  std::string data;
  std::vector<std::pair<std::string, std::string>> kv;
  {
    std::string header;  // num_shards=1, endianness little (default, omitted), version{producer=1}
    put_tag(&header, 1, 0);
    put_varint(&header, 1);
    std::string ver;
    put_tag(&ver, 1, 0);
    put_varint(&ver, 1);
    put_bytes(&header, 3, ver);
    kv.emplace_back("", header);
  }
  for (auto& t : tensors_) {
    const Pending& p = t.second;
    std::string e;
    put_tag(&e, 1, 0);
    put_varint(&e, static_cast<uint64_t>(p.dtype));
    std::string shape;
    for (int64_t d : p.shape) {
      std::string dim;
      put_tag(&dim, 1, 0);
      put_varint(&dim, static_cast<uint64_t>(d));
      put_bytes(&shape, 2, dim);
    }
    put_bytes(&e, 2, shape);
    if (!data.empty()) {
      put_tag(&e, 4, 0);
      put_varint(&e, data.size());
    }
    put_tag(&e, 5, 0);
    put_varint(&e, p.bytes.size());
    uint32_t crc;
    if (p.dtype == DT_STRING) {
      // crc over (length as u32) then (stored length-crc) then payload: matches TF's writer.
      Cursor c{p.bytes.data(), p.bytes.data() + p.bytes.size()};
      uint32_t l32 = static_cast<uint32_t>(c.varint());
      uint32_t running = crc32c(reinterpret_cast<const uint8_t*>(&l32), 4);
      running = crc32c(c.p, static_cast<size_t>(c.end - c.p), running);
      crc = crc_mask(running);
    } else {
      crc = crc_mask(crc32c(p.bytes.data(), p.bytes.size()));
    }
    put_tag(&e, 6, 5);
    put_u32(&e, crc);
    kv.emplace_back(t.first, e);
    data.append(reinterpret_cast<const char*>(p.bytes.data()), p.bytes.size());
  }

  // One data block with prefix compression (restart every 16 keys), empty metaindex,
  // one-entry index block, footer.
  auto build_block = [](const std::vector<std::pair<std::string, std::string>>& items) {
    std::string b;
    std::vector<uint32_t> restarts;
    std::string prev;
    for (size_t i = 0; i < items.size(); ++i) {
      size_t shared = 0;
      if (i % 16 == 0) {
        restarts.push_back(static_cast<uint32_t>(b.size()));
      } else {
        size_t lim = std::min(prev.size(), items[i].first.size());
        while (shared < lim && prev[shared] == items[i].first[shared]) ++shared;
      }
      put_varint(&b, shared);
      put_varint(&b, items[i].first.size() - shared);
      put_varint(&b, items[i].second.size());
      b.append(items[i].first, shared, std::string::npos);
      b.append(items[i].second);
      prev = items[i].first;
    }
    if (restarts.empty()) restarts.push_back(0);
    for (uint32_t r : restarts) put_u32(&b, r);
    put_u32(&b, static_cast<uint32_t>(restarts.size()));
    return b;
  };
  auto append_block = [](std::string* file, const std::string& block) {
    BlockHandle h{file->size(), block.size()};
    file->append(block);
    file->push_back('\0');
    uint32_t crc = crc_mask(crc32c(reinterpret_cast<const uint8_t*>(file->data()) + h.offset, block.size() + 1));
    put_u32(file, crc);
    return h;
  };
  std::string file;
  BlockHandle data_h = append_block(&file, build_block(kv));
  BlockHandle meta_h = append_block(&file, build_block({}));
  std::string handle;
  put_varint(&handle, data_h.offset);
  put_varint(&handle, data_h.size);
  std::string last_key = kv.back().first;
  BlockHandle index_h = append_block(&file, build_block({{last_key, handle}}));
  std::string footer;
  put_varint(&footer, meta_h.offset);
  put_varint(&footer, meta_h.size);
  put_varint(&footer, index_h.offset);
  put_varint(&footer, index_h.size);
  footer.resize(40, '\0');
  uint64_t magic = kTableMagic;
  footer.append(reinterpret_cast<const char*>(&magic), 8);
  file.append(footer);

  auto dump = [](const std::string& path, const std::string& bytes) {
    std::ofstream f(path, std::ios::binary | std::ios::trunc);
    if (!f || !f.write(bytes.data(), static_cast<std::streamsize>(bytes.size()))) fail("cannot write " + path);
  };
  dump(prefix + ".index", file);
  dump(prefix + ".data-00000-of-00001", data);
}

}  // namespace gnnqc
