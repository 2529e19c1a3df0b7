// model_loader.cpp -- OBJ / MTL reader of the host side (reference: src/asset_utils/model_loader.cpp:20-365).
//
// Produces, for any input the reference accepts, the same CpuGeometry, the same material table (including its
// order, which is the iteration order of a std::unordered_map filled in file order -- model_loader.cpp:289-294,
// SURVEY.md section 9 Q23) and therefore the same de-indexed vertex buffer, triangle list and BVH as the reference.
// Reference quirks that are kept because they shape the uploaded arrays (Q6, Q24):
//   * faces that come before the first `usemtl` join the first named group; faces of a file with no `usemtl`
//     at all are dropped (model_loader.cpp:144-153,172-174)
//   * a repeated `newmtl` name does not open a new material and does not stop the following property lines
//     from overwriting the material that was current before it (model_loader.cpp:200-219)
//   * has_texcoords is never raised, so every uploaded uv is (0,0) (model_loader.cpp:322-324)
// Lines are tokenised by hand (strtof / strtol) instead of one istringstream per line and per face corner.
#include "asset_utils/model_loader.h"

#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>

namespace AssetUtils {
namespace {

std::string g_object_folder = "./objects/";

const char* kBlank = " \n\r\t";

// A line cut into whitespace-separated words with stream-like number extraction: once an extraction fails,
// every later one on the same line fails too (what `linestream >> a >> b >> c` does).
class Words {
 public:
  explicit Words(const std::string& line) : p_(line.c_str()) {}

  bool Next(std::string* out) {
    Skip();
    if (!ok_ || *p_ == '\0') return ok_ = false;
    const char* b = p_;
    while (*p_ != '\0' && std::strchr(kBlank, *p_) == nullptr) ++p_;
    out->assign(b, p_);
    return true;
  }
  bool Next(float* out) {
    Skip();
    if (!ok_) return false;
    char* end = nullptr;
    const float v = std::strtof(p_, &end);
    if (end == p_) return ok_ = false;
    p_ = end;
    *out = v;
    return true;
  }

 private:
  void Skip() {
    while (*p_ != '\0' && std::strchr(kBlank, *p_) != nullptr) ++p_;
  }
  const char* p_;
  bool ok_ = true;
};

bool Trimmed(std::string* line) {
  const std::size_t b = line->find_first_not_of(kBlank);
  if (b == std::string::npos) return false;
  const std::size_t e = line->find_last_not_of(kBlank);
  *line = line->substr(b, e - b + 1);
  return (*line)[0] != '#';
}

// "12", "12/7", "12//3", "12/7/3" -> zero-based indices appended to the lists that are present
void ParseCorner(const std::string& word, std::vector<std::uint32_t>* v, std::vector<std::uint32_t>* vt,
                 std::vector<std::uint32_t>* vn) {
  const std::size_t s1 = word.find('/');
  const std::size_t s2 = s1 == std::string::npos ? std::string::npos : word.find('/', s1 + 1);
  const std::string part[3] = {
      word.substr(0, s1),
      s1 == std::string::npos ? std::string() : word.substr(s1 + 1, s2 == std::string::npos ? s2 : s2 - s1 - 1),
      s2 == std::string::npos ? std::string() : word.substr(s2 + 1)};
  std::vector<std::uint32_t>* dst[3] = {v, vt, vn};
  for (int k = 0; k < 3; ++k) {
    if (part[k].empty()) continue;
    const long one_based = std::stol(part[k]);  // throws on garbage, as the reference does
    dst[k]->push_back(static_cast<std::uint32_t>(one_based - 1));
  }
}

}  // namespace

void SetObjectFolder(const std::string& folder) {
  g_object_folder = folder;
  if (!g_object_folder.empty() && g_object_folder.back() != '/') g_object_folder += '/';
}
const std::string& GetObjectFolder() { return g_object_folder; }

std::unique_ptr<Model> LoadObject(const std::string& name) {
  const std::string folder = g_object_folder + name + "/";
  std::vector<std::string> mtl_files;
  auto geometry = Detail::ParseOBJ(folder + name + ".obj", &mtl_files);

  std::unordered_map<std::string, Material> materials;
  for (const std::string& mtl : mtl_files) Detail::ParseMTL(folder, mtl, &materials);

  if (!geometry) throw std::runtime_error("error getting geo");
  return Detail::ConvertCPUGeometryToModel(std::move(geometry), std::move(materials));
}

namespace Detail {

std::unique_ptr<CpuGeometry> ParseOBJ(const std::string& file_path, std::vector<std::string>* mtl_files) {
  std::ifstream file(file_path);
  if (!file) {
    std::cerr << "Error: Cannot open file " << file_path << std::endl;
    return nullptr;
  }
  auto geo = std::make_unique<CpuGeometry>();
  CpuSubGeometry group;  // faces collected since the last `usemtl`
  std::vector<std::uint32_t> v, vt, vn;
  std::string line, key, word;

  while (std::getline(file, line)) {
    if (!Trimmed(&line)) continue;
    Words words(line);
    words.Next(&key);

    if (key == "v" || key == "vn") {
      glm::vec3 p;
      if (key == "vn") geo->has_normals = true;
      if (words.Next(&p.x) && words.Next(&p.y) && words.Next(&p.z))
        (key == "v" ? geo->vertices : geo->normals).push_back(p);
      else
        std::cerr << (key == "v" ? "Failed to read 3 floats for vertex." : "Failed to read 3 floats for normal.")
                  << std::endl;
    } else if (key == "vt") {
      glm::vec2 uv;
      if (words.Next(&uv.x) && words.Next(&uv.y)) {
        geo->textures.push_back(uv);
        // the reference never raises has_texcoords (Q6: every uploaded uv is zero); the texture option does
        if (TexturesEnabled()) geo->has_texcoords = true;
      }
      else
        std::cerr << "Failed to read 2 floats for texcoord." << std::endl;
    } else if (key == "f") {
      v.clear();
      vt.clear();
      vn.clear();
      while (words.Next(&word)) ParseCorner(word, &v, &vt, &vn);
      if (v.size() != 3 && v.size() != 4) {
        std::cerr << "Unexpected face vertex count: " << v.size() << std::endl;
        continue;
      }
      // triangle (0,1,2); a quad adds (0,2,3)
      Face first(v[0], v[1], v[2]);
      if (vn.size() > 2) {
        first.normal_idxs = {vn[0], vn[1], vn[2]};
        first.SetNormalIdxsValid();
      }
      if (vt.size() > 2) {
        first.texture_idxs = {vt[0], vt[1], vt[2]};
        first.SetTextureIdxsValid();
      }
      group.faces.push_back(first);
      if (v.size() == 4) {
        Face second(v[0], v[2], v[3]);
        if (vn.size() == 4) {
          second.normal_idxs = {vn[0], vn[2], vn[3]};
          second.SetNormalIdxsValid();
        }
        if (vt.size() == 4) {
          second.texture_idxs = {vt[0], vt[2], vt[3]};
          second.SetTextureIdxsValid();
        }
        group.faces.push_back(second);
      }
    } else if (key == "usemtl") {
      if (!group.material.empty()) {
        geo->geometries.push_back(std::move(group));
        group = CpuSubGeometry();
      }
      std::string name;
      words.Next(&name);
      group.material = name;
    } else if (key == "mtllib") {
      std::string name;
      words.Next(&name);
      mtl_files->push_back(name);
    } else if (key == "s" || key == "o" || key == "g") {
      // smoothing groups, object and group names carry nothing the tracer uses
    } else {
      std::cout << "unhandled obj line prefix: " << key << std::endl;
    }
  }
  if (!group.material.empty()) geo->geometries.push_back(std::move(group));
  return geo;
}

void ParseMTL(const std::string& folder_path, const std::string& file_name,
              std::unordered_map<std::string, Material>* libs) {
  std::ifstream file(folder_path + file_name);
  if (!file) {
    std::cerr << "Error: Cannot open file " << folder_path + file_name << std::endl;
    return;
  }
  Material* current = nullptr;
  std::string line, key;
  while (std::getline(file, line)) {
    if (!Trimmed(&line)) continue;
    Words words(line);
    words.Next(&key);

    if (key == "newmtl") {
      std::string name;
      words.Next(&name);
      if (libs->count(name) != 0)
        std::cout << "multiple materials with the same name, skipping" << std::endl;  // `current` stays (Q24)
      else
        current = &libs->emplace(name, Material{}).first->second;
      continue;
    }
    if (current == nullptr) {
      std::cerr << "Material not defined" << std::endl;
      continue;
    }
    if (key == "Kd" || key == "Ks") {
      glm::vec3 c(0.0f);  // a short line leaves zeros (a failed stream extraction stores 0)
      if (words.Next(&c.x) && words.Next(&c.y)) words.Next(&c.z);
      (key == "Kd" ? current->diffuse : current->specular) = c;
    } else if (key == "Ns") {
      float e = 0.0f;
      words.Next(&e);
      current->specular_ex = e;
    } else if (key == "map_Kd") {
      std::string texture_name;
      words.Next(&texture_name);
      current->use_texture = true;
      current->texture = GPUTexture(folder_path + "/" + texture_name, true);
    } else if (key == "Ka" || key == "Tf" || key == "Ni" || key == "Ke" || key == "d" || key == "illum" ||
               key == "Tr" || key == "map_Ka") {
      // ambient, transmission, refraction index, emission, dissolve, illumination model: unused by the tracer
    } else {
      std::cout << "Unknown material property: " << key << std::endl;
    }
  }
}

std::unique_ptr<Model> ConvertCPUGeometryToModel(std::unique_ptr<CpuGeometry> cpu_geo,
                                                 std::unordered_map<std::string, Material> materials) {
  // material slot = position in the map's iteration order
  std::vector<Material> table;
  table.reserve(materials.size());
  std::unordered_map<std::string, std::uint32_t> slot_of;
  slot_of.reserve(materials.size());
  for (const auto& entry : materials) {
    slot_of[entry.first] = static_cast<std::uint32_t>(table.size());
    table.push_back(entry.second);
  }

  std::size_t n_faces = 0;
  for (const CpuSubGeometry& g : cpu_geo->geometries) n_faces += g.faces.size();
  std::vector<GPU::PackedVertexData> corners;  // three fresh vertices per triangle, nothing shared
  corners.reserve(3 * n_faces);
  std::vector<GPU::Triangle> triangles;
  triangles.reserve(n_faces);

  for (const CpuSubGeometry& g : cpu_geo->geometries) {
    std::uint32_t slot = 0;
    const auto found = slot_of.find(g.material);
    if (found != slot_of.end())
      slot = found->second;
    else
      std::cout << "Couldn't find expected material" << std::endl;

    for (const Face& face : g.faces) {
      GPU::Triangle tri;
      tri.material_idx = slot;
      for (int k = 0; k < 3; ++k) {
        glm::vec2 uv(0.0f);
        if (cpu_geo->has_texcoords && face.IsTextureIdxsValid()) uv = cpu_geo->textures[face.texture_idxs[k]];
        tri.vertex_idxs[k] = static_cast<std::uint32_t>(corners.size());
        corners.emplace_back(cpu_geo->vertices[face.vertex_idxs[k]], uv);
      }
      triangles.push_back(tri);
    }
  }

  const auto centre = [&corners](const GPU::Triangle& t) -> glm::vec3 {
    return (corners[t.vertex_idxs[0]].vertex + corners[t.vertex_idxs[1]].vertex + corners[t.vertex_idxs[2]].vertex) /
           3.0f;
  };
  const auto bounds = [&corners](const GPU::Triangle& t) -> std::pair<glm::vec3, glm::vec3> {
    const glm::vec3& a = corners[t.vertex_idxs[0]].vertex;
    const glm::vec3& b = corners[t.vertex_idxs[1]].vertex;
    const glm::vec3& c = corners[t.vertex_idxs[2]].vertex;
    return {glm::min(glm::min(a, b), c), glm::max(glm::max(a, b), c)};
  };
  IntersectionUtils::BVH<GPU::Triangle> bvh(std::move(triangles), centre, bounds);
  return std::make_unique<Model>(std::move(bvh), std::move(table), std::move(corners));
}

}  // namespace Detail
}  // namespace AssetUtils
