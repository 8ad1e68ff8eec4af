// Model.hpp — same public interface as the reference's Model (src/Model.hpp:8-54): loads a glTF file into per-submesh
// vertex / index arrays, the material -> image index table and RGBA8 images.
#pragma once

#include "Math.hpp"

#include <cstdint>
#include <string>
#include <vector>

class Model final
{
public:
    struct Vertex
    {
        vkrt::vec4 position{};
        vkrt::vec4 normal{};
        vkrt::vec4 uv{};
        vkrt::vec4 tangent{};
    };

    struct Material
    {
        int baseColor = -1;
        int metallicRoughnessImage = -1;
        int normalImage = -1;
    };

    struct Image
    {
        unsigned int width;
        unsigned int height;
        unsigned int components;
        unsigned int bitsPerChannel;
        std::vector<unsigned char> data;
    };

    using Index = uint32_t;

    struct Submesh
    {
        std::vector<Vertex> vertices;
        std::vector<Index> indices;
        int material = -1;
    };

    // `filename` is relative to the models folder (MODELS_FOLDER at build time, overridable at run time with the
    // VKRT_MODELS_FOLDER environment variable or setModelsFolder()).
    Model(const std::string& filename);
    ~Model() {}

    static void setModelsFolder(const std::string& folder);
    static const std::string& modelsFolder();

    std::vector<Submesh> submeshes;
    std::vector<Material> materials;
    std::vector<Image> images;

    uint64_t vertexBufferSizeInBytes = 0;
    uint64_t indexBufferSizeInBytes = 0;
};
