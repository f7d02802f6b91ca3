// HDF5 subset reader/writer for Keras ".h5" checkpoints (there is no libhdf5 / h5py on the box).
//
// What Keras 2.x + h5py (libver "earliest") writes, and therefore what is implemented:
//   superblock v0; "old style" groups (symbol-table message -> v1 B-tree + local heap + SNOD);
//   v1 object headers with continuation blocks; contiguous (and compact) dataset layout;
//   datatypes: little-endian fixed-point, IEEE float, fixed-length strings, variable-length strings
//   (read only, through the global heap); attributes stored as header messages.
// Not implemented (clear error instead of garbage): chunked/compressed datasets, dense attribute
// storage, v2 object headers, big-endian types.
// The writer emits the same subset with large B-tree fan-out (leaf K 16, internal K 64) so every
// group fits a single-level B-tree (<= 4096 children).
//
// Reference binding: model.load_weights(by_name, skip_mismatch) / ModelCheckpoint / load_model at
// train.py:77,113-114,406.  Format facts follow the public HDF5 file
// format specification v1.1/2.0; no code was taken from libhdf5.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/b200_rt.h"

namespace h5 {

constexpr uint64_t UNDEF = ~0ull;
thread_local std::string g_error;

struct TypeInfo {
  int cls = -1;        // 0 float, 1 int, 2 fixed string, 3 vlen string
  int size = 0;        // element size in the file
  bool is_signed = true;
};

struct Attr {
  std::string name;
  TypeInfo type;
  std::vector<long long> dims;
  std::vector<uint8_t> data;  // fixed-size elements; vlen strings are resolved to '\0'-joined text
  std::vector<std::string> strings;  // for vlen strings
};

struct Node {
  bool is_group = true;
  std::map<std::string, std::unique_ptr<Node>> children;
  std::vector<std::string> child_order;
  std::vector<Attr> attrs;
  // dataset
  TypeInfo type;
  std::vector<long long> dims;
  std::vector<uint8_t> data;      // writer: owned copy; reader: filled lazily
  uint64_t data_addr = UNDEF, data_size = 0;
  bool loaded = false;
  uint64_t header_addr = 0;
};

inline void put(std::vector<uint8_t>& b, const void* p, size_t n) {
  const uint8_t* c = static_cast<const uint8_t*>(p);
  b.insert(b.end(), c, c + n);
}
inline void put8(std::vector<uint8_t>& b, uint8_t v) { b.push_back(v); }
inline void put16(std::vector<uint8_t>& b, uint16_t v) { put(b, &v, 2); }
inline void put32(std::vector<uint8_t>& b, uint32_t v) { put(b, &v, 4); }
inline void put64(std::vector<uint8_t>& b, uint64_t v) { put(b, &v, 8); }
inline void pad8(std::vector<uint8_t>& b) {
  while (b.size() % 8) b.push_back(0);
}

// ------------------------------------------------------------------------------------------ writer
std::vector<uint8_t> encode_datatype(const TypeInfo& t) {
  std::vector<uint8_t> m;
  if (t.cls == 1) {
    put8(m, 0x10 | 0);                       // version 1, class 0 fixed-point
    put8(m, t.is_signed ? 0x08 : 0x00);      // little-endian, signed bit
    put8(m, 0);
    put8(m, 0);
    put32(m, t.size);
    put16(m, 0);
    put16(m, static_cast<uint16_t>(t.size * 8));
  } else if (t.cls == 0) {
    put8(m, 0x10 | 1);                       // class 1 floating-point
    put8(m, 0x20);                           // LE, mantissa normalisation: msb implied
    put8(m, static_cast<uint8_t>(t.size * 8 - 1));  // sign bit position
    put8(m, 0);
    put32(m, t.size);
    put16(m, 0);
    put16(m, static_cast<uint16_t>(t.size * 8));
    if (t.size == 4) {
      put8(m, 23); put8(m, 8); put8(m, 0); put8(m, 23); put32(m, 127);
    } else {
      put8(m, 52); put8(m, 11); put8(m, 0); put8(m, 52); put32(m, 1023);
    }
  } else {
    put8(m, 0x10 | 3);                       // class 3 string
    put8(m, 0x01);                           // null-padded, ASCII (what h5py uses for numpy 'S')
    put8(m, 0);
    put8(m, 0);
    put32(m, t.size);
  }
  pad8(m);
  return m;
}

std::vector<uint8_t> encode_dataspace(const std::vector<long long>& dims) {
  std::vector<uint8_t> m;
  put8(m, 1);
  put8(m, static_cast<uint8_t>(dims.size()));
  put8(m, 0);
  put8(m, 0);
  put32(m, 0);
  for (long long d : dims) put64(m, static_cast<uint64_t>(d));
  return m;
}

void add_message(std::vector<uint8_t>& body, uint16_t type, const std::vector<uint8_t>& data, int& count) {
  std::vector<uint8_t> d = data;
  pad8(d);
  put16(body, type);
  put16(body, static_cast<uint16_t>(d.size()));
  put8(body, 0);
  put8(body, 0); put8(body, 0); put8(body, 0);
  put(body, d.data(), d.size());
  ++count;
}

// Scalar byte strings are stored the way h5py stores a Python `bytes` / `str` attribute (Keras' keras_version, backend,
// model_config, training_config): a VARIABLE-LENGTH string whose characters live in a global heap collection and whose
// attribute message holds a 16-byte reference.  That is also the only way a 100 KB model_config fits: a version-1
// header message cannot exceed 64 KiB.
struct VlenRef {
  uint64_t heap_addr = 0;
  uint32_t index = 0;
};

bool is_vlen_scalar(const Attr& a) { return a.type.cls == 2 && a.dims.empty(); }

std::vector<uint8_t> encode_vlen_string_type() {
  std::vector<uint8_t> m;
  put8(m, 0x10 | 9);   // version 1, class 9 variable-length
  put8(m, 0x01);       // type: string; padding: null-terminated
  put8(m, 0x00);       // character set: ASCII
  put8(m, 0);
  put32(m, 16);        // size of the in-file descriptor: length, heap address, object index
  put8(m, 0x10 | 3);   // base type: 1-byte null-terminated ASCII string
  put8(m, 0);
  put8(m, 0);
  put8(m, 0);
  put32(m, 1);
  return m;
}

std::vector<uint8_t> encode_attr(const Attr& a, const VlenRef* ref = nullptr) {
  std::vector<uint8_t> dt = ref ? encode_vlen_string_type() : encode_datatype(a.type), ds = encode_dataspace(a.dims), m;
  put8(m, 1);
  put8(m, 0);
  put16(m, static_cast<uint16_t>(a.name.size() + 1));
  put16(m, static_cast<uint16_t>(dt.size()));
  put16(m, static_cast<uint16_t>(ds.size()));
  put(m, a.name.c_str(), a.name.size() + 1);
  pad8(m);
  put(m, dt.data(), dt.size());
  pad8(m);
  put(m, ds.data(), ds.size());
  pad8(m);
  if (ref) {
    put32(m, static_cast<uint32_t>(a.data.size()));
    put64(m, ref->heap_addr);
    put32(m, ref->index);
  } else {
    put(m, a.data.data(), a.data.size());
  }
  return m;
}

struct Writer {
  std::vector<uint8_t> file;
  static constexpr int kLeafK = 16, kInternalK = 64;

  uint64_t alloc(size_t n) {
    while (file.size() % 8) file.push_back(0);
    const uint64_t a = file.size();
    file.resize(file.size() + n, 0);
    return a;
  }
  void write_at(uint64_t addr, const std::vector<uint8_t>& b) { memcpy(file.data() + addr, b.data(), b.size()); }

  uint64_t write_header(const std::vector<uint8_t>& body, int nmsg) {
    std::vector<uint8_t> h;
    put8(h, 1);
    put8(h, 0);
    put16(h, static_cast<uint16_t>(nmsg));
    put32(h, 1);
    put32(h, static_cast<uint32_t>(body.size()));
    put32(h, 0);
    put(h, body.data(), body.size());
    const uint64_t a = alloc(h.size());
    write_at(a, h);
    return a;
  }

  // ---- global heap: one collection with every variable-length string of the file, allocated before the objects
  std::map<const Attr*, VlenRef> vlen;

  void collect_vlen(const Node& n, std::vector<const Attr*>& out) {
    for (const Attr& a : n.attrs)
      if (is_vlen_scalar(a)) out.push_back(&a);
    for (auto& kv : n.children) collect_vlen(*kv.second, out);
  }

  void write_global_heap(const Node& root) {
    std::vector<const Attr*> all;
    collect_vlen(root, all);
    if (all.empty()) return;
    size_t need = 16 + 16;                       // collection header + the free-space object
    for (const Attr* a : all) need += 16 + ((a->data.size() + 7) & ~size_t(7));
    const size_t size = need < 4096 ? 4096 : ((need + 4095) & ~size_t(4095));     // libhdf5's minimum collection size
    std::vector<uint8_t> g;
    put(g, "GCOL", 4);
    put8(g, 1); put8(g, 0); put8(g, 0); put8(g, 0);
    put64(g, size);
    const uint64_t addr = alloc(size);
    uint16_t idx = 1;
    for (const Attr* a : all) {
      put16(g, idx);
      put16(g, 1);                               // reference count
      put32(g, 0);
      put64(g, a->data.size());
      put(g, a->data.data(), a->data.size());
      pad8(g);
      vlen[a] = VlenRef{addr, idx};
      ++idx;
    }
    const size_t free_bytes = size - g.size();
    put16(g, 0);                                 // object 0: the free space, its size counts its own header
    put16(g, 0);
    put32(g, 0);
    put64(g, free_bytes);
    write_at(addr, g);
  }

  std::vector<uint8_t> attr_message(const Attr& a) {
    auto it = vlen.find(&a);
    return encode_attr(a, it == vlen.end() ? nullptr : &it->second);
  }

  bool attrs_fit(const Node& n, std::string& why) {
    for (const Attr& a : n.attrs)
      if (attr_message(a).size() + 8 > 65000) {
        why = "attribute '" + a.name + "' exceeds the 64 KiB object-header message limit (split it, as Keras does for layer_names)";
        return false;
      }
    return true;
  }

  uint64_t write_node(Node& n) {
    std::string why;
    if (!attrs_fit(n, why)) {
      g_error = why;
      return UNDEF;
    }
    if (!n.is_group) {
      const uint64_t daddr = n.data.empty() ? UNDEF : alloc(n.data.size());
      if (!n.data.empty()) memcpy(file.data() + daddr, n.data.data(), n.data.size());
      std::vector<uint8_t> body;
      int cnt = 0;
      add_message(body, 0x0001, encode_dataspace(n.dims), cnt);
      add_message(body, 0x0003, encode_datatype(n.type), cnt);
      std::vector<uint8_t> fill = {2, 1, 2, 0};  // v2, early allocation, write fill if set, undefined fill
      add_message(body, 0x0005, fill, cnt);
      std::vector<uint8_t> lay;
      put8(lay, 3);
      put8(lay, 1);
      put64(lay, daddr);
      put64(lay, n.data.size());
      add_message(body, 0x0008, lay, cnt);
      for (const Attr& a : n.attrs) add_message(body, 0x000C, attr_message(a), cnt);
      return n.header_addr = write_header(body, cnt);
    }
    // children first
    std::vector<std::pair<std::string, uint64_t>> kids;
    for (auto& kv : n.children) {
      const uint64_t a = write_node(*kv.second);
      if (a == UNDEF) return UNDEF;
      kids.emplace_back(kv.first, a);
    }
    std::sort(kids.begin(), kids.end());
    if (kids.size() > static_cast<size_t>(2 * kLeafK) * 2 * kInternalK) {
      g_error = "group has too many children for the single-level B-tree this writer emits";
      return UNDEF;
    }
    // local heap data segment: "" at 0, then names
    std::vector<uint8_t> seg(8, 0);
    std::vector<uint64_t> name_off;
    for (auto& k : kids) {
      name_off.push_back(seg.size());
      put(seg, k.first.c_str(), k.first.size() + 1);
      pad8(seg);
    }
    const uint64_t free_off = seg.size();
    put64(seg, 1);    // next free block: none (H5HL_FREE_NULL)
    put64(seg, 16);   // size of this free block
    const uint64_t seg_addr = alloc(seg.size());
    write_at(seg_addr, seg);
    std::vector<uint8_t> heap;
    put(heap, "HEAP", 4);
    put8(heap, 0); put8(heap, 0); put8(heap, 0); put8(heap, 0);
    put64(heap, seg.size());
    put64(heap, free_off);
    put64(heap, seg_addr);
    const uint64_t heap_addr = alloc(heap.size());
    write_at(heap_addr, heap);
    // symbol nodes
    const size_t per = 2 * kLeafK;
    std::vector<uint64_t> snods, last_name;
    for (size_t i = 0; i < kids.size(); i += per) {
      const size_t n_here = std::min(per, kids.size() - i);
      std::vector<uint8_t> s;
      put(s, "SNOD", 4);
      put8(s, 1);
      put8(s, 0);
      put16(s, static_cast<uint16_t>(n_here));
      for (size_t j = 0; j < per; ++j) {
        if (j < n_here) {
          put64(s, name_off[i + j]);
          put64(s, kids[i + j].second);
          put32(s, 0);
          put32(s, 0);
          put64(s, 0); put64(s, 0);
        } else {
          for (int z = 0; z < 5; ++z) put64(s, 0);
        }
      }
      const uint64_t a = alloc(s.size());
      write_at(a, s);
      snods.push_back(a);
      last_name.push_back(name_off[i + n_here - 1]);
    }
    std::vector<uint8_t> bt;
    put(bt, "TREE", 4);
    put8(bt, 0);
    put8(bt, 0);
    put16(bt, static_cast<uint16_t>(snods.size()));
    put64(bt, UNDEF);
    put64(bt, UNDEF);
    put64(bt, 0);  // key 0: the empty string
    for (int i = 0; i < 2 * kInternalK; ++i) {
      if (i < static_cast<int>(snods.size())) {
        put64(bt, snods[i]);
        put64(bt, last_name[i]);
      } else {
        put64(bt, 0);
        put64(bt, 0);
      }
    }
    const uint64_t bt_addr = alloc(bt.size());
    write_at(bt_addr, bt);
    std::vector<uint8_t> body;
    int cnt = 0;
    std::vector<uint8_t> st;
    put64(st, bt_addr);
    put64(st, heap_addr);
    add_message(body, 0x0011, st, cnt);
    for (const Attr& a : n.attrs) add_message(body, 0x000C, attr_message(a), cnt);
    n.header_addr = write_header(body, cnt);
    btree_of[n.header_addr] = {bt_addr, heap_addr};
    return n.header_addr;
  }

  std::map<uint64_t, std::pair<uint64_t, uint64_t>> btree_of;

  bool write_file(Node& root, const std::string& path) {
    file.assign(96, 0);
    write_global_heap(root);
    const uint64_t root_addr = write_node(root);
    if (root_addr == UNDEF) return false;
    while (file.size() % 8) file.push_back(0);
    std::vector<uint8_t> sb;
    const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    put(sb, sig, 8);
    put8(sb, 0); put8(sb, 0); put8(sb, 0); put8(sb, 0); put8(sb, 0);
    put8(sb, 8); put8(sb, 8); put8(sb, 0);
    put16(sb, kLeafK);
    put16(sb, kInternalK);
    put32(sb, 0);
    put64(sb, 0);
    put64(sb, UNDEF);
    put64(sb, file.size());
    put64(sb, UNDEF);
    put64(sb, 0);            // root entry: link name offset
    put64(sb, root_addr);
    put32(sb, 1);            // cache type 1: B-tree and heap addresses cached in the scratch pad
    put32(sb, 0);
    put64(sb, btree_of[root_addr].first);
    put64(sb, btree_of[root_addr].second);
    write_at(0, sb);
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) {
      g_error = "cannot open " + path + " for writing";
      return false;
    }
    const size_t w = fwrite(file.data(), 1, file.size(), f);
    fclose(f);
    if (w != file.size()) {
      g_error = "short write to " + path;
      return false;
    }
    return true;
  }
};

}  // namespace h5

#include "hdf5_reader.inc"
