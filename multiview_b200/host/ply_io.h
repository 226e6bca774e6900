// ply_io.h -- reader of the triangle mesh given with --mesh (rig_calibrator.cc:1298-1299 loadMeshBuildTree, which loads a .ply through
// MVE's mve::geom::load_ply_mesh).  Only what the ray / mesh intersection needs is kept: float32 vertex positions and the vertex
// indices of the faces, the arrays mvba_mesh_intersect takes.  Formats: ascii and binary_little_endian 1.0; vertex properties
// other than x, y, z (normals, colours, confidences) and elements other than vertex / face are skipped; polygons with more than
// three corners are split into a fan.  Errors are std::runtime_error.
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace mvba_host {

struct TriangleMesh {
  std::vector<float> vertices;   // 3 per vertex
  std::vector<uint32_t> faces;   // 3 per triangle
};

namespace ply_detail {
inline int type_size(const std::string& t) {
  if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
  if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
  if (t == "int" || t == "uint" || t == "float" || t == "int32" || t == "uint32" || t == "float32") return 4;
  if (t == "double" || t == "float64") return 8;
  throw std::runtime_error("ply: unknown property type " + t);
}
inline double read_binary(std::istream& in, const std::string& t) {
  char b[8];
  const int n = type_size(t);
  in.read(b, n);
  if (!in) throw std::runtime_error("ply: unexpected end of file");
  if (t == "char" || t == "int8") { int8_t v; std::memcpy(&v, b, 1); return v; }
  if (t == "uchar" || t == "uint8") { uint8_t v; std::memcpy(&v, b, 1); return v; }
  if (t == "short" || t == "int16") { int16_t v; std::memcpy(&v, b, 2); return v; }
  if (t == "ushort" || t == "uint16") { uint16_t v; std::memcpy(&v, b, 2); return v; }
  if (t == "int" || t == "int32") { int32_t v; std::memcpy(&v, b, 4); return v; }
  if (t == "uint" || t == "uint32") { uint32_t v; std::memcpy(&v, b, 4); return v; }
  if (t == "float" || t == "float32") { float v; std::memcpy(&v, b, 4); return v; }
  double v; std::memcpy(&v, b, 8); return v;
}
struct Property { std::string name, type, count_type; bool is_list = false; };
struct Element { std::string name; int64_t count = 0; std::vector<Property> props; };
}  // namespace ply_detail

inline TriangleMesh read_ply(const std::string& path) {
  using namespace ply_detail;
  std::ifstream in(path, std::ios::binary);
  if (!in) throw std::runtime_error("Cannot open mesh file: " + path);
  std::string line;
  std::getline(in, line);
  if (line.substr(0, 3) != "ply") throw std::runtime_error("Not a ply file: " + path);
  bool ascii = true, have_format = false;
  std::vector<Element> elements;
  while (std::getline(in, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    std::istringstream ls(line);
    std::string key;
    ls >> key;
    if (key == "end_header") break;
    if (key == "format") {
      std::string f; ls >> f;
      if (f == "ascii") ascii = true;
      else if (f == "binary_little_endian") ascii = false;
      else throw std::runtime_error("ply: unsupported format " + f);
      have_format = true;
    } else if (key == "element") {
      Element e; ls >> e.name >> e.count; elements.push_back(e);
    } else if (key == "property") {
      if (elements.empty()) throw std::runtime_error("ply: property before element");
      Property p; std::string t; ls >> t;
      if (t == "list") { p.is_list = true; ls >> p.count_type >> p.type >> p.name; }
      else { p.type = t; ls >> p.name; }
      type_size(p.type);
      elements.back().props.push_back(p);
    }
  }
  if (!have_format) throw std::runtime_error("ply: no format line in " + path);
  TriangleMesh mesh;
  auto next = [&](const std::string& type) -> double {
    if (!ascii) return read_binary(in, type);
    double v; if (!(in >> v)) throw std::runtime_error("ply: unexpected end of file"); return v;
  };
  for (const Element& e : elements) {
    for (int64_t k = 0; k < e.count; k++) {
      float xyz[3] = {0, 0, 0};
      for (const Property& p : e.props) {
        if (!p.is_list) {
          const double v = next(p.type);
          if (e.name == "vertex") { if (p.name == "x") xyz[0] = (float)v; else if (p.name == "y") xyz[1] = (float)v; else if (p.name == "z") xyz[2] = (float)v; }
          continue;
        }
        const int n = (int)next(p.count_type);
        std::vector<uint32_t> idx((size_t)n);
        for (int j = 0; j < n; j++) idx[j] = (uint32_t)next(p.type);
        if (e.name == "face" && (p.name == "vertex_indices" || p.name == "vertex_index"))
          for (int j = 1; j + 1 < n; j++) { mesh.faces.push_back(idx[0]); mesh.faces.push_back(idx[j]); mesh.faces.push_back(idx[j + 1]); }
      }
      if (e.name == "vertex") mesh.vertices.insert(mesh.vertices.end(), xyz, xyz + 3);
    }
  }
  const size_t nv = mesh.vertices.size() / 3;
  for (uint32_t i : mesh.faces) if (i >= nv) throw std::runtime_error("ply: face index out of range in " + path);
  return mesh;
}

}  // namespace mvba_host
