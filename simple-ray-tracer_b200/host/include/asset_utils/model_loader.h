// asset_utils/model_loader.h -- OBJ/MTL loading, API of the reference's include/asset_utils/model_loader.h:15-23.
#pragma once

#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "asset_utils/types.h"

namespace AssetUtils {

// Loads <objects folder>/<name>/<name>.obj with its material libraries and builds the model's BVH.
// Throws std::runtime_error("error getting geo") when the OBJ cannot be opened (model_loader.cpp:28-29).
std::unique_ptr<Model> LoadObject(const std::string& name);

// The reference hard-codes "./objects/" (model_loader.cpp:14); that stays the default.
void SetObjectFolder(const std::string& folder);
const std::string& GetObjectFolder();

namespace Detail {
std::unique_ptr<CpuGeometry> ParseOBJ(const std::string& file_path, std::vector<std::string>* mtl_files);
void ParseMTL(const std::string& folder_path, const std::string& file_name,
              std::unordered_map<std::string, Material>* libs);
std::unique_ptr<Model> ConvertCPUGeometryToModel(std::unique_ptr<CpuGeometry> cpu_geo,
                                                 std::unordered_map<std::string, Material> materials);
}  // namespace Detail
}  // namespace AssetUtils
