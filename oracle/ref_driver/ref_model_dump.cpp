// ref_model_dump.cpp — drives the REFERENCE's own Model loader (src/Model.cpp + vendored tinygltf / stb_image, compiled
// in place from /root/reference) on sponza/Sponza.gltf and prints FNV-1a checksums of everything it produces.
// Built only into oracle/_ref/ by oracle/build_ref.sh; its output is committed as tests/golden/model_golden.json and
// the host Model (vkrt_b200/host/Model.cpp) is checked against it.  Optional argv[1]: file to receive raw image bytes.
#include "Model.hpp"

#include <cstdint>
#include <cstdio>

static uint64_t fnv(const void* data, size_t n, uint64_t h = 1469598103934665603ull)
{
    const unsigned char* p = static_cast<const unsigned char*>(data);
    for (size_t i = 0; i < n; ++i)
    {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    return h;
}

int main(int argc, char** argv)
{
    Model model("sponza/Sponza.gltf");
    FILE* raw = argc > 1 ? fopen(argv[1], "wb") : nullptr;
    printf("{\n  \"vertexBufferSizeInBytes\": %llu,\n  \"indexBufferSizeInBytes\": %llu,\n", (unsigned long long)model.vertexBufferSizeInBytes,
           (unsigned long long)model.indexBufferSizeInBytes);
    printf("  \"submeshes\": [\n");
    for (size_t i = 0; i < model.submeshes.size(); ++i)
    {
        const Model::Submesh& s = model.submeshes[i];
        printf("    {\"vertices\": %zu, \"indices\": %zu, \"material\": %d, \"vertexHash\": \"%016llx\", \"indexHash\": \"%016llx\"}%s\n",
               s.vertices.size(), s.indices.size(), s.material, (unsigned long long)fnv(s.vertices.data(), s.vertices.size() * sizeof(Model::Vertex)),
               (unsigned long long)fnv(s.indices.data(), s.indices.size() * sizeof(Model::Index)), i + 1 < model.submeshes.size() ? "," : "");
    }
    printf("  ],\n  \"materials\": [");
    for (size_t i = 0; i < model.materials.size(); ++i)
    {
        printf("[%d, %d, %d]%s", model.materials[i].baseColor, model.materials[i].metallicRoughnessImage, model.materials[i].normalImage,
               i + 1 < model.materials.size() ? ", " : "");
    }
    printf("],\n  \"images\": [\n");
    for (size_t i = 0; i < model.images.size(); ++i)
    {
        const Model::Image& im = model.images[i];
        printf("    {\"width\": %u, \"height\": %u, \"components\": %u, \"bits\": %u, \"hash\": \"%016llx\"}%s\n", im.width, im.height, im.components,
               im.bitsPerChannel, (unsigned long long)fnv(im.data.data(), im.data.size()), i + 1 < model.images.size() ? "," : "");
        if (raw)
        {
            fwrite(im.data.data(), 1, im.data.size(), raw);
        }
    }
    printf("  ]\n}\n");
    if (raw)
    {
        fclose(raw);
    }
    return 0;
}
