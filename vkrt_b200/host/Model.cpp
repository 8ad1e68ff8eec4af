// Model.cpp — glTF loader with the reference loader's semantics (src/Model.cpp:48-191), without tinygltf/stb:
//   * only meshes[0] is read, node transforms are ignored (:50-53);
//   * indices are always read as uint16 and the read pointer advances by bufferView.byteStride + element size (:68-76);
//   * vertex attributes advance by bufferView.byteStride, no tight-packing fallback (:115); POSITION / NORMAL /
//     TEXCOORD_0 / TANGENT are copied into the vec4 slots of Model::Vertex, other attributes only size the array;
//   * materials keep only the image indices of base colour, metallic-roughness and normal textures (:122-136);
//   * images are decoded to 8-bit RGBA (tinygltf forces 4 components, submodules/tinygltf/tiny_gltf.h:2363);
//   * any load problem aborts, like the reference's CHECK / LOGE (:165-178).
// Images are decoded in parallel on std::thread workers (the reference decodes 69 images on one thread).
#include "Model.hpp"
#include "ImageDecode.hpp"
#include "Json.hpp"
#include "Utils.hpp"

#include <atomic>
#include <cstring>
#include <fstream>
#include <thread>

#ifndef MODELS_FOLDER
#define MODELS_FOLDER "./models/"
#endif

namespace
{
std::string g_modelsFolder;

bool readFile(const std::string& path, std::vector<uint8_t>& out)
{
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f)
    {
        return false;
    }
    const std::streamsize n = f.tellg();
    f.seekg(0);
    out.resize((size_t)n);
    return n == 0 || (bool)f.read(reinterpret_cast<char*>(out.data()), n);
}

size_t componentTypeSize(long long componentType)
{
    switch (componentType)
    {
    case 5120: case 5121: return 1; // BYTE, UNSIGNED_BYTE
    case 5122: case 5123: return 2; // SHORT, UNSIGNED_SHORT
    case 5124: case 5125: case 5126: return 4; // INT, UNSIGNED_INT, FLOAT
    case 5130: return 8; // DOUBLE
    default: return 0;
    }
}

size_t typeCount(const std::string& type)
{
    if (type == "SCALAR") return 1;
    if (type == "VEC2") return 2;
    if (type == "VEC3") return 3;
    if (type == "VEC4") return 4;
    return 0;
}

size_t accessorElementSize(const vkrt::JsonValue& accessor)
{
    const size_t c = componentTypeSize(accessor["componentType"].asInt());
    const size_t t = typeCount(accessor["type"].asString());
    CHECK(c != 0 && t != 0);
    return c * t;
}

int sourceOrMinusOne(const vkrt::JsonValue& textures, const vkrt::JsonValue& textureInfo)
{
    if (!textureInfo.has("index"))
    {
        return -1;
    }
    const long long index = textureInfo["index"].asInt(-1);
    if (index < 0)
    {
        return -1;
    }
    return (int)textures[(size_t)index]["source"].asInt(-1);
}
} // namespace

void Model::setModelsFolder(const std::string& folder)
{
    g_modelsFolder = folder;
    if (!g_modelsFolder.empty() && g_modelsFolder.back() != '/')
    {
        g_modelsFolder += '/';
    }
}

const std::string& Model::modelsFolder()
{
    if (g_modelsFolder.empty())
    {
        const char* env = std::getenv("VKRT_MODELS_FOLDER");
        setModelsFolder(env ? env : MODELS_FOLDER);
    }
    return g_modelsFolder;
}

Model::Model(const std::string& filename)
{
    const std::string filepath = modelsFolder() + filename;
    const std::string baseDir = filepath.substr(0, filepath.find_last_of('/') + 1);
    printf("Loading model %s... ", filepath.c_str());

    std::vector<uint8_t> text;
    if (!readFile(filepath, text))
    {
        LOGE(("File not found : " + filepath).c_str());
    }
    vkrt::JsonValue gltf;
    std::string error;
    if (!vkrt::JsonParser::parse(std::string(text.begin(), text.end()), gltf, error))
    {
        LOGE(error.c_str());
    }
    CHECK(gltf["meshes"].size() != 0);

    // buffers
    std::vector<std::vector<uint8_t>> buffers(gltf["buffers"].size());
    for (size_t i = 0; i < buffers.size(); ++i)
    {
        const std::string uri = gltf["buffers"][i]["uri"].asString();
        if (!readFile(baseDir + uri, buffers[i]))
        {
            LOGE(("File not found : " + uri).c_str());
        }
        CHECK(buffers[i].size() >= (size_t)gltf["buffers"][i]["byteLength"].asInt());
    }

    // submeshes (src/Model.cpp:48-120)
    const vkrt::JsonValue& primitives = gltf["meshes"][0]["primitives"];
    const vkrt::JsonValue& accessors = gltf["accessors"];
    const vkrt::JsonValue& bufferViews = gltf["bufferViews"];
    submeshes.resize(primitives.size());
    for (size_t i = 0; i < primitives.size(); ++i)
    {
        const vkrt::JsonValue& primitive = primitives[i];
        submeshes[i].material = (int)primitive["material"].asInt(-1);
        {
            const vkrt::JsonValue& accessor = accessors[(size_t)primitive["indices"].asInt()];
            const vkrt::JsonValue& bufferView = bufferViews[(size_t)accessor["bufferView"].asInt()];
            const std::vector<uint8_t>& buffer = buffers[(size_t)bufferView["buffer"].asInt()];
            std::vector<uint32_t>& indices = submeshes[i].indices;
            const size_t count = (size_t)accessor["count"].asInt();
            indices.resize(count);
            const size_t elementSize = accessorElementSize(accessor);
            const size_t offset = (size_t)bufferView["byteOffset"].asInt() + (size_t)accessor["byteOffset"].asInt();
            const size_t stride = (size_t)bufferView["byteStride"].asInt() + elementSize;
            const size_t lastIndex = offset + (size_t)bufferView["byteLength"].asInt() - 1;
            size_t p = offset;
            for (size_t k = 0; k < count; ++k)
            {
                CHECK(p < lastIndex && p + 2 <= buffer.size());
                unsigned short value = 0;
                std::memcpy(&value, &buffer[p], sizeof(unsigned short)); // always uint16, as the reference
                indices[k] = value;
                p += stride;
            }
        }
        // std::map iteration order == tinygltf's attribute map order (sorted by name)
        for (const auto& attribute : primitive["attributes"].object)
        {
            const std::string& name = attribute.first;
            const vkrt::JsonValue& accessor = accessors[(size_t)attribute.second.asInt()];
            const vkrt::JsonValue& bufferView = bufferViews[(size_t)accessor["bufferView"].asInt()];
            const std::vector<uint8_t>& buffer = buffers[(size_t)bufferView["buffer"].asInt()];
            std::vector<Vertex>& vertices = submeshes[i].vertices;
            const size_t count = (size_t)accessor["count"].asInt();
            vertices.resize(count);
            const size_t elementSize = accessorElementSize(accessor);
            const size_t offset = (size_t)bufferView["byteOffset"].asInt() + (size_t)accessor["byteOffset"].asInt();
            const size_t stride = (size_t)bufferView["byteStride"].asInt();
            const size_t lastIndex = offset + (size_t)bufferView["byteLength"].asInt() - 1;
            size_t p = offset;
            for (size_t k = 0; k < count; ++k)
            {
                CHECK(p < lastIndex && p + elementSize <= buffer.size());
                if (name == "POSITION")
                {
                    std::memcpy(&vertices[k].position, &buffer[p], std::min<size_t>(elementSize, 16));
                }
                else if (name == "NORMAL")
                {
                    std::memcpy(&vertices[k].normal, &buffer[p], std::min<size_t>(elementSize, 16));
                }
                else if (name == "TEXCOORD_0")
                {
                    std::memcpy(&vertices[k].uv, &buffer[p], std::min<size_t>(elementSize, 16));
                }
                else if (name == "TANGENT")
                {
                    std::memcpy(&vertices[k].tangent, &buffer[p], std::min<size_t>(elementSize, 16));
                }
                p += stride;
            }
        }
    }

    // materials (src/Model.cpp:122-136)
    const vkrt::JsonValue& textures = gltf["textures"];
    materials.resize(gltf["materials"].size());
    for (size_t i = 0; i < materials.size(); ++i)
    {
        const vkrt::JsonValue& m = gltf["materials"][i];
        const vkrt::JsonValue& pbr = m["pbrMetallicRoughness"];
        materials[i].baseColor = sourceOrMinusOne(textures, pbr["baseColorTexture"]);
        materials[i].metallicRoughnessImage = sourceOrMinusOne(textures, pbr["metallicRoughnessTexture"]);
        materials[i].normalImage = sourceOrMinusOne(textures, m["normalTexture"]);
    }

    // images (src/Model.cpp:138-151), decoded in parallel
    images.resize(gltf["images"].size());
    std::vector<std::string> errors(images.size());
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        for (;;)
        {
            const size_t i = next.fetch_add(1);
            if (i >= images.size())
            {
                break;
            }
            const std::string uri = gltf["images"][i]["uri"].asString();
            std::vector<uint8_t> file;
            if (!readFile(baseDir + uri, file))
            {
                errors[i] = "File not found : " + uri;
                continue;
            }
            vkrt::DecodedImage decoded;
            if (!vkrt::decodeImage(file.data(), file.size(), decoded, errors[i]))
            {
                errors[i] = uri + ": " + errors[i];
                continue;
            }
            images[i].width = decoded.width;
            images[i].height = decoded.height;
            images[i].components = 4;
            images[i].bitsPerChannel = 8;
            images[i].data = std::move(decoded.rgba);
        }
    };
    const unsigned threads = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)images.size()));
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t)
    {
        pool.emplace_back(worker);
    }
    worker();
    for (std::thread& t : pool)
    {
        t.join();
    }
    for (const std::string& e : errors)
    {
        if (!e.empty())
        {
            LOGE(e.c_str());
        }
    }

    for (const Model::Submesh& submesh : submeshes)
    {
        vertexBufferSizeInBytes += sizeof(Model::Vertex) * submesh.vertices.size();
        indexBufferSizeInBytes += sizeof(Model::Index) * submesh.indices.size();
    }
    printf("Completed\n");
}
