// textures.cu — the mip chain the reference builds for every texture with a chain of linear blits
// (Raytracer::createMipmaps, src/Raytracer.cpp:572-640, level count :467-468).  The reference's ray-tracing shaders never
// read it (texture() without derivatives samples level 0); here it feeds the optional ray-cone level-of-detail
// mode of the shading kernels (vkrt_trace_set_texture_lod).
//
// One launch per level over ALL images (blockIdx.y = image), each thread one destination texel = blitTexel() of the
// previous level.  HBM stream: level k reads 4 B and writes 1 B per source texel, 475 MB in total for the 68 1024x1024
// images of Sponza; the levels of one image are packed one after the other, images 256-byte aligned in one pool.
#include "context.hpp"
#include "shading.cuh"

#include <algorithm>

namespace vkrt
{

namespace
{
struct MipJob
{
    const uint8_t* src;
    uint8_t* dst; // nullptr: this image has no such level
    uint32_t srcW, srcH, dstW, dstH;
};

__global__ void __launch_bounds__(256) k_mip_level(const MipJob* __restrict__ jobs)
{
    // c / 255.0f for every byte value, divided once per block: the 16 IEEE divisions per texel made the first version
    // instruction-bound (1.1 TB/s of DRAM traffic on level 1); the table holds the identical quotients
    __shared__ float unorm[256];
    unorm[threadIdx.x] = (float)threadIdx.x / 255.0f;
    __syncthreads();
    const MipJob job = jobs[blockIdx.y];
    if (!job.dst)
    {
        return;
    }
    const uint32_t total = job.dstW * job.dstH;
    uint32_t* out = reinterpret_cast<uint32_t*>(job.dst);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
    {
        out[i] = blitTexel(job.src, job.srcW, job.srcH, job.dstW, job.dstH, i % job.dstW, i / job.dstW, unorm);
    }
}
} // namespace

int generateMips(Context* ctx)
{
    const uint32_t count = ctx->textureCount;
    cudaStream_t s = ctx->stream;
    ctx->mipsReady = false;
    std::vector<TextureMips> mips(count);
    std::vector<size_t> offsets(count);
    size_t total = 0;
    uint32_t maxLevels = 1;
    for (uint32_t i = 0; i < count; ++i)
    {
        const TextureRecord& t = ctx->textureHost[i];
        const uint32_t levels = mipLevelCount(t.width, t.height);
        maxLevels = std::max(maxLevels, levels);
        mips[i].levels = levels, mips[i].pad = 0;
        offsets[i] = total;
        size_t bytes = 0;
        for (uint32_t k = 1; k < levels; ++k)
        {
            bytes += (size_t)mipExtent(t.width, k) * mipExtent(t.height, k) * 4;
        }
        total += (bytes + 255) & ~(size_t)255;
    }
    VKRT_CUDA_TRY(ctx, ctx->mipPool.reserve(std::max<size_t>(total, 256)));
    uint8_t* pool = ctx->mipPool.as<uint8_t>();

    // job table: level-major, one row of `count` jobs per level
    std::vector<MipJob> jobs((size_t)(maxLevels - 1) * count);
    std::vector<uint32_t> levelMaxTexels(maxLevels, 0);
    for (uint32_t i = 0; i < count; ++i)
    {
        const TextureRecord& t = ctx->textureHost[i];
        mips[i].texels = pool + offsets[i];
        const uint8_t* src = t.texels;
        uint8_t* dst = pool + offsets[i];
        for (uint32_t k = 1; k < maxLevels; ++k)
        {
            MipJob& j = jobs[(size_t)(k - 1) * count + i];
            if (k >= mips[i].levels)
            {
                j = MipJob{nullptr, nullptr, 0, 0, 0, 0};
                continue;
            }
            j.src = src, j.dst = dst;
            j.srcW = mipExtent(t.width, k - 1), j.srcH = mipExtent(t.height, k - 1);
            j.dstW = mipExtent(t.width, k), j.dstH = mipExtent(t.height, k);
            levelMaxTexels[k] = std::max(levelMaxTexels[k], j.dstW * j.dstH);
            src = dst;
            dst += (size_t)j.dstW * j.dstH * 4;
        }
    }
    if (!jobs.empty())
    {
        DeviceBuffer table;
        VKRT_CUDA_TRY(ctx, table.reserve(jobs.size() * sizeof(MipJob)));
        cudaError_t e = cudaMemcpyAsync(table.ptr, jobs.data(), jobs.size() * sizeof(MipJob), cudaMemcpyHostToDevice, s);
        for (uint32_t k = 1; k < maxLevels && e == cudaSuccess; ++k)
        {
            // enough blocks per image to fill the machine together with the other images, at most 4 texels per thread
            const uint32_t want = (levelMaxTexels[k] + 1023u) / 1024u;
            const uint32_t cap = std::max<uint32_t>(1u, (uint32_t)(ctx->smCount * 8) / std::max(count, 1u)); // one resident wave
            const dim3 grid(std::max(1u, std::min(want, cap)), count);
            k_mip_level<<<grid, 256, 0, s>>>(table.as<MipJob>() + (size_t)(k - 1) * count);
            ctx->launches += 1;
            e = cudaGetLastError();
        }
        if (e == cudaSuccess)
        {
            e = cudaStreamSynchronize(s);
        }
        table.release();
        if (e != cudaSuccess)
        {
            return ctx->fail(VKRT_ERROR_CUDA, std::string("vkrt_textures_generate_mips: ") + cudaGetErrorString(e));
        }
    }
    VKRT_CUDA_TRY(ctx, ctx->textureMips.reserve(std::max<size_t>(mips.size() * sizeof(TextureMips), 16)));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->textureMips.ptr, mips.data(), mips.size() * sizeof(TextureMips), cudaMemcpyHostToDevice, s));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    ctx->mipsHost = mips;
    ctx->mipsReady = true;
    return VKRT_SUCCESS;
}

} // namespace vkrt
