// Weight loading for the LeNet classifier, replacing Caffe's
// Net::CopyTrainedLayersFrom (reference caffe_classifier.cpp:13-14).
// A .caffemodel is a serialized caffe.NetParameter; only the protobuf WIRE FORMAT
// is parsed here (no protobuf dependency):
//   NetParameter : layers = 2 (V1LayerParameter), layer = 100 (LayerParameter)
//   LayerParameter   : name = 1, blobs = 7
//   V1LayerParameter : name = 4, blobs = 6
//   BlobProto        : data = 5 (packed or repeated float), shape = 7, num/channels/height/width = 1..4
// Layers are matched by the names in caffe/sigmoid_lenet.prototxt: conv1, conv2, ip1, ip2.
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "ag2_internal.h"

namespace ag2 {
namespace {

struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  bool ok = true;
  bool more() const { return ok && p < end; }
  uint64_t varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      if (p >= end) {
        ok = false;
        return 0;
      }
      uint8_t b = *p++;
      v |= (uint64_t)(b & 0x7f) << shift;
      if (!(b & 0x80)) return v;
    }
    ok = false;
    return 0;
  }
  Reader sub() {
    uint64_t len = varint();
    Reader r{p, p, ok};
    if (!ok || len > (uint64_t)(end - p)) {
      ok = false;
      r.ok = false;
      return r;
    }
    r.end = p + len;
    p += len;
    return r;
  }
  void skip(int wire) {
    switch (wire) {
      case 0: varint(); break;
      case 1: if (end - p < 8) ok = false; else p += 8; break;
      case 2: sub(); break;
      case 5: if (end - p < 4) ok = false; else p += 4; break;
      default: ok = false;
    }
  }
};

bool parse_blob(Reader r, std::vector<float>& data) {
  data.clear();
  while (r.more()) {
    uint64_t tag = r.varint();
    int field = (int)(tag >> 3), wire = (int)(tag & 7);
    if (field == 5 && wire == 2) {  // packed floats
      Reader d = r.sub();
      if (!r.ok) return false;
      size_t n = (size_t)(d.end - d.p) / 4;
      size_t old = data.size();
      data.resize(old + n);
      memcpy(data.data() + old, d.p, n * 4);
    } else if (field == 5 && wire == 5) {
      if (r.end - r.p < 4) return false;
      float f;
      memcpy(&f, r.p, 4);
      r.p += 4;
      data.push_back(f);
    } else {
      r.skip(wire);
    }
  }
  return r.ok;
}

// one layer message; name_field / blobs_field differ between V1 and V2
bool parse_layer(Reader r, int name_field, int blobs_field, std::string& name,
                 std::vector<std::vector<float>>& blobs) {
  name.clear();
  blobs.clear();
  while (r.more()) {
    uint64_t tag = r.varint();
    int field = (int)(tag >> 3), wire = (int)(tag & 7);
    if (field == name_field && wire == 2) {
      Reader s = r.sub();
      if (!r.ok) return false;
      name.assign((const char*)s.p, (size_t)(s.end - s.p));
    } else if (field == blobs_field && wire == 2) {
      Reader b = r.sub();
      if (!r.ok) return false;
      blobs.emplace_back();
      if (!parse_blob(b, blobs.back())) return false;
    } else {
      r.skip(wire);
    }
  }
  return r.ok;
}

}  // namespace

int load_caffemodel(const char* path, std::vector<float> out[8], std::string& err) {
  FILE* f = fopen(path, "rb");
  if (!f) {
    err = std::string("cannot open ") + path;
    return AG2_E_IO;
  }
  std::vector<uint8_t> buf;
  uint8_t tmp[1 << 16];
  size_t got;
  while ((got = fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + got);
  fclose(f);
  static const char* names[4] = {"conv1", "conv2", "ip1", "ip2"};
  static const size_t wsize[4] = {20 * 75, 50 * 500, 500 * 7200, 2 * 500};
  static const size_t bsize[4] = {20, 50, 500, 2};
  bool found[4] = {false, false, false, false};
  Reader r{buf.data(), buf.data() + buf.size(), true};
  while (r.more()) {
    uint64_t tag = r.varint();
    int field = (int)(tag >> 3), wire = (int)(tag & 7);
    if ((field == 100 || field == 2) && wire == 2) {
      Reader l = r.sub();
      if (!r.ok) break;
      std::string name;
      std::vector<std::vector<float>> blobs;
      if (!parse_layer(l, field == 100 ? 1 : 4, field == 100 ? 7 : 6, name, blobs)) {
        err = "malformed layer message";
        return AG2_E_IO;
      }
      for (int k = 0; k < 4; k++) {
        if (name != names[k] || blobs.empty()) continue;
        if (blobs[0].size() != wsize[k] || (blobs.size() > 1 && blobs[1].size() != bsize[k])) {
          err = "layer " + name + ": blob sizes do not match caffe/sigmoid_lenet.prototxt";
          return AG2_E_IO;
        }
        out[2 * k] = blobs[0];
        out[2 * k + 1] = blobs.size() > 1 ? blobs[1] : std::vector<float>(bsize[k], 0.f);
        found[k] = true;
      }
    } else {
      r.skip(wire);
    }
  }
  if (!r.ok) {
    err = "malformed caffemodel (wire format)";
    return AG2_E_IO;
  }
  for (int k = 0; k < 4; k++)
    if (!found[k]) {
      err = std::string("layer ") + names[k] + " with weights not found";
      return AG2_E_IO;
    }
  return AG2_OK;
}

}  // namespace ag2

// The same parser without a context: the blobs of a .caffemodel into host arrays (NULL = not wanted).
extern "C" int ag2_read_lenet_caffemodel(const char* path, float* conv1_w, float* conv1_b, float* conv2_w,
                                         float* conv2_b, float* ip1_w, float* ip1_b, float* ip2_w, float* ip2_b) {
  if (!path) return AG2_E_INVALID;
  std::vector<float> w[8];
  std::string err;
  int rc = ag2::load_caffemodel(path, w, err);
  if (rc) return rc;
  float* dst[8] = {conv1_w, conv1_b, conv2_w, conv2_b, ip1_w, ip1_b, ip2_w, ip2_b};
  for (int i = 0; i < 8; i++)
    if (dst[i]) memcpy(dst[i], w[i].data(), w[i].size() * sizeof(float));
  return AG2_OK;
}

namespace ag2 {

// Caffe XavierFiller: uniform(-s, s), s = sqrt(3 / fan_in), fan_in = count / num_output;
// bias filler defaults to constant 0.  Deterministic splitmix64 stream.
void xavier_fill(uint64_t seed, std::vector<float> out[8]) {
  static const size_t wsize[4] = {20 * 75, 50 * 500, 500 * 7200, 2 * 500};
  static const size_t nout[4] = {20, 50, 500, 2};
  uint64_t state = seed;
  auto next = [&state]() {
    state += 0x9E3779B97F4A7C15ull;
    uint64_t x = state;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
  };
  for (int k = 0; k < 4; k++) {
    double fan_in = (double)wsize[k] / (double)nout[k];
    double s = sqrt(3.0 / fan_in);
    out[2 * k].resize(wsize[k]);
    for (size_t i = 0; i < wsize[k]; i++) {
      double u = (double)(next() >> 11) * (1.0 / 9007199254740992.0);  // [0,1)
      out[2 * k][i] = (float)((2.0 * u - 1.0) * s);
    }
    out[2 * k + 1].assign(nout[k], 0.f);
  }
}

}  // namespace ag2
