// f3: on-disk / wire formats around the path (host C++, no device work).
//  * PCD reader replacing pcl::io::loadPCDFile<PointXYZRGBA / PointXYZRGBNormal>
//    (reference cloud_camera.cpp:231-240): DATA ascii | binary | binary_compressed (LZF,
//    fields stored one after the other), fields x y z [rgb | rgba] [normal_x normal_y
//    normal_z] [anything else, skipped], SIZE / TYPE / COUNT honoured.
//  * PCD writer (ascii / binary) for tools and tests.
//  * GraspListMsg serialisation: the ROS 1 wire format of agile_grasp2/GraspListMsg
//    (msg/GraspListMsg.msg, msg/GraspMsg.msg, GraspHypothesis::convertToGraspMsg
//    grasp_hypothesis.cpp:37-50, createGraspListMsg grasp_detection_node.cpp:306-313):
//    Header {uint32 seq, uint32 secs, uint32 nsecs, string frame_id} + uint32 count +
//    per grasp 6 x 3 float64 (surface bottom top axis approach binormal) + float32 width +
//    float32 score, little endian.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <sstream>
#include <string>
#include <vector>

#include "../../include/ag2.h"

namespace {

struct Field {
  std::string name;
  int size = 4;
  char type = 'F';
  int count = 1;
  size_t offset = 0;  // byte offset inside a point record
};

struct PcdHeader {
  std::vector<Field> fields;
  size_t points = 0, width = 0, height = 1;
  int data = -1;  // 0 ascii, 1 binary, 2 binary_compressed
  size_t record = 0;
  long data_pos = 0;
};

bool parse_header(FILE* f, PcdHeader& h) {
  char line[4096];
  bool have_points = false;
  while (fgets(line, sizeof(line), f)) {
    if (line[0] == '#') continue;
    std::istringstream is(line);
    std::string key;
    is >> key;
    if (key == "VERSION" || key == "VIEWPOINT" || key.empty()) continue;
    if (key == "FIELDS" || key == "COLUMNS") {
      std::string s;
      while (is >> s) {
        Field fd;
        fd.name = s;
        h.fields.push_back(fd);
      }
    } else if (key == "SIZE") {
      for (Field& fd : h.fields) is >> fd.size;
    } else if (key == "TYPE") {
      for (Field& fd : h.fields) is >> fd.type;
    } else if (key == "COUNT") {
      for (Field& fd : h.fields) is >> fd.count;
    } else if (key == "WIDTH") {
      is >> h.width;
    } else if (key == "HEIGHT") {
      is >> h.height;
    } else if (key == "POINTS") {
      is >> h.points;
      have_points = true;
    } else if (key == "DATA") {
      std::string s;
      is >> s;
      h.data = s == "ascii" ? 0 : (s == "binary" ? 1 : (s == "binary_compressed" ? 2 : -1));
      h.data_pos = ftell(f);
      break;
    }
  }
  if (!have_points) h.points = h.width * h.height;
  size_t off = 0;
  for (Field& fd : h.fields) {
    fd.offset = off;
    off += (size_t)fd.size * (size_t)fd.count;
  }
  h.record = off;
  return h.data >= 0 && !h.fields.empty() && h.record > 0;
}

double read_scalar(const unsigned char* p, const Field& fd) {
  switch (fd.type) {
    case 'F':
      if (fd.size == 4) { float v; memcpy(&v, p, 4); return v; }
      if (fd.size == 8) { double v; memcpy(&v, p, 8); return v; }
      break;
    case 'U':
      if (fd.size == 1) return *p;
      if (fd.size == 2) { uint16_t v; memcpy(&v, p, 2); return v; }
      if (fd.size == 4) { uint32_t v; memcpy(&v, p, 4); return v; }
      break;
    case 'I':
      if (fd.size == 1) return *(const int8_t*)p;
      if (fd.size == 2) { int16_t v; memcpy(&v, p, 2); return v; }
      if (fd.size == 4) { int32_t v; memcpy(&v, p, 4); return v; }
      break;
  }
  return NAN;
}

// liblzf decompressor (the format PCL uses for binary_compressed)
bool lzf_decompress(const unsigned char* in, size_t in_len, unsigned char* out, size_t out_len) {
  size_t ip = 0, op = 0;
  while (ip < in_len) {
    unsigned ctrl = in[ip++];
    if (ctrl < 32) {  // literal run
      size_t len = ctrl + 1;
      if (op + len > out_len || ip + len > in_len) return false;
      memcpy(out + op, in + ip, len);
      op += len, ip += len;
    } else {  // back reference
      size_t len = ctrl >> 5;
      if (len == 7) {
        if (ip >= in_len) return false;
        len += in[ip++];
      }
      if (ip >= in_len) return false;
      size_t ref_off = ((size_t)(ctrl & 0x1f) << 8) + in[ip++] + 1;
      len += 2;
      if (ref_off > op || op + len > out_len) return false;
      size_t ref = op - ref_off;
      for (size_t k = 0; k < len; k++) out[op++] = out[ref++];
    }
  }
  return op == out_len;
}

int find_field(const PcdHeader& h, const char* name) {
  for (size_t i = 0; i < h.fields.size(); i++)
    if (h.fields[i].name == name) return (int)i;
  return -1;
}

void put_u32(std::vector<uint8_t>& b, uint32_t v) {
  for (int k = 0; k < 4; k++) b.push_back((uint8_t)(v >> (8 * k)));
}
void put_f64(std::vector<uint8_t>& b, double v) {
  uint64_t u;
  memcpy(&u, &v, 8);
  for (int k = 0; k < 8; k++) b.push_back((uint8_t)(u >> (8 * k)));
}
void put_f32(std::vector<uint8_t>& b, float v) {
  uint32_t u;
  memcpy(&u, &v, 4);
  put_u32(b, u);
}

}  // namespace

extern "C" {

int ag2_pcd_info(const char* path, size_t* n, int* has_normals, int* has_color) {
  if (!path || !n) return AG2_E_INVALID;
  FILE* f = fopen(path, "rb");
  if (!f) return AG2_E_IO;
  PcdHeader h;
  bool ok = parse_header(f, h);
  fclose(f);
  if (!ok || find_field(h, "x") < 0 || find_field(h, "y") < 0 || find_field(h, "z") < 0) return AG2_E_IO;
  *n = h.points;
  if (has_normals)
    *has_normals = find_field(h, "normal_x") >= 0 && find_field(h, "normal_y") >= 0 && find_field(h, "normal_z") >= 0;
  if (has_color) *has_color = find_field(h, "rgba") >= 0 || find_field(h, "rgb") >= 0;
  return AG2_OK;
}

int ag2_pcd_read(const char* path, float* xyz, float* normals_or_null, uint32_t* rgba_or_null, size_t cap,
                 size_t* n_out) {
  if (!path || !xyz) return AG2_E_INVALID;
  FILE* f = fopen(path, "rb");
  if (!f) return AG2_E_IO;
  PcdHeader h;
  if (!parse_header(f, h)) {
    fclose(f);
    return AG2_E_IO;
  }
  const int fx = find_field(h, "x"), fy = find_field(h, "y"), fz = find_field(h, "z");
  const int fn[3] = {find_field(h, "normal_x"), find_field(h, "normal_y"), find_field(h, "normal_z")};
  int fc = find_field(h, "rgba");
  if (fc < 0) fc = find_field(h, "rgb");
  if (fx < 0 || fy < 0 || fz < 0) {
    fclose(f);
    return AG2_E_IO;
  }
  const size_t n = h.points;
  if (n_out) *n_out = n;
  if (n > cap) {
    fclose(f);
    return AG2_E_CAPACITY;
  }
  const bool want_n = normals_or_null && fn[0] >= 0 && fn[1] >= 0 && fn[2] >= 0;
  int rc = AG2_OK;
  if (h.data == 0) {  // ascii: one point per line, fields in order
    std::vector<char> line(1 << 16);
    for (size_t i = 0; i < n && rc == AG2_OK; i++) {
      if (!fgets(line.data(), (int)line.size(), f)) {
        rc = AG2_E_IO;
        break;
      }
      char* p = line.data();
      for (size_t k = 0; k < h.fields.size(); k++) {
        const Field& fd = h.fields[k];
        for (int c = 0; c < fd.count; c++) {
          char* end = nullptr;
          double v;
          uint32_t bits = 0;
          if (fd.type == 'F') {
            v = strtod(p, &end);
            float fv = (float)v;
            memcpy(&bits, &fv, 4);
          } else {
            unsigned long long u = strtoull(p, &end, 10);
            v = (double)u;
            bits = (uint32_t)u;
          }
          if (end == p) {
            rc = AG2_E_IO;
            break;
          }
          p = end;
          if (c) continue;
          if ((int)k == fx) xyz[3 * i] = (float)v;
          if ((int)k == fy) xyz[3 * i + 1] = (float)v;
          if ((int)k == fz) xyz[3 * i + 2] = (float)v;
          if (want_n)
            for (int a = 0; a < 3; a++)
              if ((int)k == fn[a]) normals_or_null[3 * i + a] = (float)v;
          if (rgba_or_null && (int)k == fc) rgba_or_null[i] = bits;  // rgb as float carries the packed bits
        }
      }
    }
  } else {
    std::vector<unsigned char> raw(n * h.record);
    bool soa = false;
    if (h.data == 1) {
      if (fread(raw.data(), 1, raw.size(), f) != raw.size()) rc = AG2_E_IO;
    } else {
      uint32_t comp = 0, uncomp = 0;
      if (fread(&comp, 4, 1, f) != 1 || fread(&uncomp, 4, 1, f) != 1 || uncomp != raw.size()) {
        rc = AG2_E_IO;
      } else {
        std::vector<unsigned char> cbuf(comp);
        if (fread(cbuf.data(), 1, comp, f) != comp || !lzf_decompress(cbuf.data(), comp, raw.data(), raw.size()))
          rc = AG2_E_IO;
      }
      soa = true;
    }
    if (rc == AG2_OK) {
      auto at = [&](size_t i, int k) -> const unsigned char* {
        const Field& fd = h.fields[k];
        // binary: array of records; binary_compressed: all of field 0, then all of field 1, ...
        return soa ? raw.data() + fd.offset * n + i * (size_t)fd.size * (size_t)fd.count
                   : raw.data() + i * h.record + fd.offset;
      };
      for (size_t i = 0; i < n; i++) {
        xyz[3 * i] = (float)read_scalar(at(i, fx), h.fields[fx]);
        xyz[3 * i + 1] = (float)read_scalar(at(i, fy), h.fields[fy]);
        xyz[3 * i + 2] = (float)read_scalar(at(i, fz), h.fields[fz]);
        if (want_n)
          for (int a = 0; a < 3; a++) normals_or_null[3 * i + a] = (float)read_scalar(at(i, fn[a]), h.fields[fn[a]]);
        if (rgba_or_null && fc >= 0) memcpy(&rgba_or_null[i], at(i, fc), 4);
      }
    }
  }
  fclose(f);
  return rc;
}

int ag2_pcd_write(const char* path, const float* xyz, const float* normals_or_null, const uint32_t* rgba_or_null,
                  size_t n, int binary) {
  if (!path || (n && !xyz)) return AG2_E_INVALID;
  FILE* f = fopen(path, "wb");
  if (!f) return AG2_E_IO;
  std::string fields = "x y z", size = "4 4 4", type = "F F F", count = "1 1 1";
  if (rgba_or_null) fields += " rgba", size += " 4", type += " U", count += " 1";
  if (normals_or_null) fields += " normal_x normal_y normal_z", size += " 4 4 4", type += " F F F", count += " 1 1 1";
  fprintf(f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS %s\nSIZE %s\nTYPE %s\nCOUNT %s\n",
          fields.c_str(), size.c_str(), type.c_str(), count.c_str());
  fprintf(f, "WIDTH %zu\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %zu\nDATA %s\n", n, n, binary ? "binary" : "ascii");
  for (size_t i = 0; i < n; i++) {
    if (binary) {
      fwrite(&xyz[3 * i], 4, 3, f);
      if (rgba_or_null) fwrite(&rgba_or_null[i], 4, 1, f);
      if (normals_or_null) fwrite(&normals_or_null[3 * i], 4, 3, f);
    } else {
      fprintf(f, "%.9g %.9g %.9g", xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      if (rgba_or_null) fprintf(f, " %u", rgba_or_null[i]);
      if (normals_or_null)
        fprintf(f, " %.9g %.9g %.9g", normals_or_null[3 * i], normals_or_null[3 * i + 1], normals_or_null[3 * i + 2]);
      fputc('\n', f);
    }
  }
  fclose(f);
  return AG2_OK;
}

int ag2_serialize_grasp_list(const ag2_hypothesis* hands, size_t n, uint32_t seq, uint32_t stamp_sec,
                             uint32_t stamp_nsec, const char* frame_id, uint8_t* out, size_t cap, size_t* len) {
  if ((n && !hands) || !len) return AG2_E_INVALID;
  std::vector<uint8_t> b;
  const size_t fl = frame_id ? strlen(frame_id) : 0;
  b.reserve(16 + fl + 4 + n * 152);
  put_u32(b, seq);
  put_u32(b, stamp_sec);
  put_u32(b, stamp_nsec);
  put_u32(b, (uint32_t)fl);
  for (size_t k = 0; k < fl; k++) b.push_back((uint8_t)frame_id[k]);
  put_u32(b, (uint32_t)n);
  for (size_t i = 0; i < n; i++) {
    const ag2_hypothesis& h = hands[i];
    const double* v[6] = {h.surface, h.bottom, h.top, h.axis, h.approach, h.binormal};
    for (int k = 0; k < 6; k++)
      for (int a = 0; a < 3; a++) put_f64(b, v[k][a]);
    put_f32(b, (float)h.width);  // std_msgs/Float32 width, score
    put_f32(b, (float)h.score);
  }
  *len = b.size();
  if (b.size() > cap) return AG2_E_CAPACITY;
  if (out) memcpy(out, b.data(), b.size());
  return AG2_OK;
}

}  // extern "C"
