#include "tfexample.h"

namespace pgn {
namespace {

struct Reader {
  const uint8_t* p; const uint8_t* end; bool ok = true;
  bool varint(uint64_t* v) {
    uint64_t r = 0; int shift = 0;
    while (p < end && shift < 64) {
      const uint8_t b = *p++;
      r |= (uint64_t)(b & 0x7f) << shift;
      if (!(b & 0x80)) { *v = r; return true; }
      shift += 7;
    }
    return ok = false;
  }
  bool bytes(std::string_view* v) {
    uint64_t n;
    if (!varint(&n) || n > (uint64_t)(end - p)) return ok = false;
    *v = std::string_view(reinterpret_cast<const char*>(p), (size_t)n);
    p += n;
    return true;
  }
  bool skip(uint32_t wire) {
    uint64_t v; std::string_view s;
    switch (wire) {
      case 0: return varint(&v);
      case 1: if (end - p < 8) return ok = false; p += 8; return true;
      case 2: return bytes(&s);
      case 5: if (end - p < 4) return ok = false; p += 4; return true;
      default: return ok = false;          // groups (3, 4) do not occur in tf.Example
    }
  }
};

Reader sub(std::string_view s) {
  Reader r{reinterpret_cast<const uint8_t*>(s.data()), reinterpret_cast<const uint8_t*>(s.data()) + s.size()};
  return r;
}

// Feature -> the single value of its BytesList, if that is what it holds
bool single_bytes(std::string_view feature, std::string_view* value) {
  Reader r = sub(feature);
  int n_values = 0; bool is_bytes = false;
  while (r.p < r.end) {
    uint64_t tag;
    if (!r.varint(&tag)) return false;
    if ((tag >> 3) == 1 && (tag & 7) == 2) {                 // bytes_list
      std::string_view list;
      if (!r.bytes(&list)) return false;
      is_bytes = true;
      Reader l = sub(list);
      while (l.p < l.end) {
        uint64_t t2;
        if (!l.varint(&t2)) return false;
        if ((t2 >> 3) == 1 && (t2 & 7) == 2) { if (!l.bytes(value)) return false; ++n_values; }
        else if (!l.skip((uint32_t)(t2 & 7))) return false;
      }
    } else if (!r.skip((uint32_t)(tag & 7))) return false;
  }
  return is_bytes && n_values == 1;
}

void put_varint(std::string* o, uint64_t v) {
  while (v >= 0x80) { o->push_back((char)(v | 0x80)); v >>= 7; }
  o->push_back((char)v);
}
size_t varint_size(uint64_t v) { size_t n = 1; while (v >= 0x80) { v >>= 7; ++n; } return n; }

}  // namespace

std::string parse_example(const uint8_t* buf, size_t len, std::vector<BytesFeature>* out) {
  out->clear();
  Reader ex{buf, buf + len};
  while (ex.p < ex.end) {
    uint64_t tag;
    if (!ex.varint(&tag)) return "truncated tf.Example";
    if ((tag >> 3) == 1 && (tag & 7) == 2) {                   // Example.features
      std::string_view features;
      if (!ex.bytes(&features)) return "truncated tf.Example (features)";
      Reader fs = sub(features);
      while (fs.p < fs.end) {
        uint64_t t2;
        if (!fs.varint(&t2)) return "truncated tf.Example (feature map)";
        if ((t2 >> 3) == 1 && (t2 & 7) == 2) {                 // one map entry
          std::string_view entry, key, feature;
          if (!fs.bytes(&entry)) return "truncated tf.Example (map entry)";
          Reader e = sub(entry);
          bool has_key = false, has_val = false;
          while (e.p < e.end) {
            uint64_t t3;
            if (!e.varint(&t3)) return "truncated tf.Example (map entry)";
            if ((t3 >> 3) == 1 && (t3 & 7) == 2) { if (!e.bytes(&key)) return "truncated tf.Example (key)"; has_key = true; }
            else if ((t3 >> 3) == 2 && (t3 & 7) == 2) { if (!e.bytes(&feature)) return "truncated tf.Example (value)"; has_val = true; }
            else if (!e.skip((uint32_t)(t3 & 7))) return "malformed tf.Example (map entry)";
          }
          std::string_view value;
          if (has_key && has_val && single_bytes(feature, &value)) {
            bool replaced = false;                             // later entries override earlier ones (protobuf maps)
            for (auto& f : *out) if (f.name == key) { f.value = value; replaced = true; }
            if (!replaced) out->push_back({key, value});
          }
        } else if (!fs.skip((uint32_t)(t2 & 7))) return "malformed tf.Example (features)";
      }
    } else if (!ex.skip((uint32_t)(tag & 7))) return "malformed tf.Example";
  }
  return "";
}

void write_example(const std::vector<BytesFeature>& feats, std::string* out) {
  out->clear();
  // sizes bottom-up: BytesList{1: value} -> Feature{1: list} -> entry{1: key, 2: feature} -> Features{1: entry...}
  std::vector<size_t> list_sz(feats.size()), feat_sz(feats.size()), entry_sz(feats.size());
  size_t features_sz = 0;
  for (size_t i = 0; i < feats.size(); ++i) {
    list_sz[i] = 1 + varint_size(feats[i].value.size()) + feats[i].value.size();
    feat_sz[i] = 1 + varint_size(list_sz[i]) + list_sz[i];
    entry_sz[i] = 1 + varint_size(feats[i].name.size()) + feats[i].name.size() + 1 + varint_size(feat_sz[i]) + feat_sz[i];
    features_sz += 1 + varint_size(entry_sz[i]) + entry_sz[i];
  }
  out->reserve(features_sz + 8);
  out->push_back((char)0x0a); put_varint(out, features_sz);                       // Example.features
  for (size_t i = 0; i < feats.size(); ++i) {
    out->push_back((char)0x0a); put_varint(out, entry_sz[i]);                     // Features.feature entry
    out->push_back((char)0x0a); put_varint(out, feats[i].name.size()); out->append(feats[i].name);      // key
    out->push_back((char)0x12); put_varint(out, feat_sz[i]);                      // value: Feature
    out->push_back((char)0x0a); put_varint(out, list_sz[i]);                      // Feature.bytes_list
    out->push_back((char)0x0a); put_varint(out, feats[i].value.size()); out->append(feats[i].value);    // BytesList.value
  }
}

}  // namespace pgn
