// comm.cu — multi-GPU tile gather over NCCL (NVLink 5 / NVSwitch).  No counterpart in the reference (single GPU).
//
// Each rank renders the 64x64 tiles it owns (tile index % size == rank) into a tile-major BGRA8 buffer that is
// directly the NCCL send buffer (written by k_resolve_sample, no staging pass).  The root receives every peer's
// buffer with one grouped ncclSend/ncclRecv exchange and scatters the tiles into the linear image.
//
// NCCL is bound at run time with dlopen so that the library loads on machines without NCCL and shares the
// already-loaded libnccl of the host process (e.g. the one bundled with PyTorch).
#include "context.hpp"

#include <dlfcn.h>
#include <mutex>

namespace vkrt
{

namespace
{
typedef struct
{
    char internal[128];
} NcclUniqueId;
typedef void* NcclComm;
enum
{
    kNcclUint8 = 1 // ncclUint8 / ncclChar family: ncclInt8=0, ncclUint8=1
};

struct NcclFns
{
    void* handle = nullptr;
    int (*getUniqueId)(NcclUniqueId*) = nullptr;
    int (*commInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*commDestroy)(NcclComm) = nullptr;
    int (*send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*broadcast)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*groupStart)() = nullptr;
    int (*groupEnd)() = nullptr;
    const char* (*getErrorString)(int) = nullptr;
    std::string error;
};

NcclFns& nccl()
{
    static NcclFns fns;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names)
        {
            fns.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (fns.handle)
            {
                break;
            }
        }
        if (!fns.handle)
        {
            fns.error = std::string("NCCL not found: ") + dlerror();
            return;
        }
        auto sym = [&](const char* s) {
            void* p = dlsym(fns.handle, s);
            if (!p && fns.error.empty())
            {
                fns.error = std::string("NCCL symbol missing: ") + s;
            }
            return p;
        };
        fns.getUniqueId = reinterpret_cast<int (*)(NcclUniqueId*)>(sym("ncclGetUniqueId"));
        fns.commInitRank = reinterpret_cast<int (*)(NcclComm*, int, NcclUniqueId, int)>(sym("ncclCommInitRank"));
        fns.commDestroy = reinterpret_cast<int (*)(NcclComm)>(sym("ncclCommDestroy"));
        fns.send = reinterpret_cast<int (*)(const void*, size_t, int, int, NcclComm, cudaStream_t)>(sym("ncclSend"));
        fns.recv = reinterpret_cast<int (*)(void*, size_t, int, int, NcclComm, cudaStream_t)>(sym("ncclRecv"));
        fns.broadcast = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t)>(sym("ncclBroadcast"));
        fns.groupStart = reinterpret_cast<int (*)()>(sym("ncclGroupStart"));
        fns.groupEnd = reinterpret_cast<int (*)()>(sym("ncclGroupEnd"));
        fns.getErrorString = reinterpret_cast<const char* (*)(int)>(sym("ncclGetErrorString"));
    });
    return fns;
}

uint32_t ownedTilesOf(uint32_t tiles, int rank, int size)
{
    return tiles > (uint32_t)rank ? (tiles - (uint32_t)rank + (uint32_t)size - 1) / (uint32_t)size : 0;
}

// Scatters the staged tiles of all peers (rank after rank, the root left out; each rank's tiles in its own order) into
// the linear image.  Rank r owns tiles r, r + size, ...: tiles / size of them, one more when r < tiles % size.
__global__ void __launch_bounds__(256)
k_untile_all(const uint32_t* __restrict__ staged, uint32_t stagedTiles, uint32_t tiles, uint32_t root, uint32_t size, uint32_t tilesX,
             uint32_t width, uint32_t height, uint32_t* __restrict__ image)
{
    const uint32_t total = stagedTiles * (uint32_t)kTilePixels;
    const uint32_t q = tiles / size, rem = tiles % size;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
    {
        uint32_t tl = i / (uint32_t)kTilePixels;
        const uint32_t pit = i % (uint32_t)kTilePixels;
        uint32_t rank = 0;
        for (; rank < size; ++rank)
        {
            const uint32_t owned = rank == root ? 0u : q + (rank < rem ? 1u : 0u);
            if (tl < owned)
            {
                break;
            }
            tl -= owned;
        }
        const uint32_t tile = rank + tl * size;
        const uint32_t x = (tile % tilesX) * (uint32_t)kTile + (pit % (uint32_t)kTile);
        const uint32_t y = (tile / tilesX) * (uint32_t)kTile + (pit / (uint32_t)kTile);
        if (rank < size && x < width && y < height)
        {
            image[(size_t)y * width + x] = staged[i];
        }
    }
}
} // namespace

int commUniqueId(uint8_t* id, std::string& err)
{
    NcclFns& n = nccl();
    if (!n.error.empty())
    {
        err = n.error;
        return VKRT_ERROR_NCCL;
    }
    NcclUniqueId uid;
    const int r = n.getUniqueId(&uid);
    if (r != 0)
    {
        err = std::string("ncclGetUniqueId: ") + n.getErrorString(r);
        return VKRT_ERROR_NCCL;
    }
    memcpy(id, uid.internal, VKRT_UNIQUE_ID_BYTES);
    return VKRT_SUCCESS;
}

int commInit(Context* ctx, const uint8_t* id, int rank, int size)
{
    NcclFns& n = nccl();
    if (!n.error.empty())
    {
        return ctx->fail(VKRT_ERROR_NCCL, n.error);
    }
    commDestroy(ctx);
    NcclUniqueId uid;
    memcpy(uid.internal, id, VKRT_UNIQUE_ID_BYTES);
    NcclComm comm = nullptr;
    const int r = n.commInitRank(&comm, size, uid, rank);
    if (r != 0)
    {
        return ctx->fail(VKRT_ERROR_NCCL, std::string("ncclCommInitRank: ") + n.getErrorString(r));
    }
    ctx->ncclComm = comm;
    ctx->rank = rank;
    ctx->size = size;
    return VKRT_SUCCESS;
}

int commDestroy(Context* ctx)
{
    if (ctx->ncclComm)
    {
        NcclFns& n = nccl();
        cudaStreamSynchronize(ctx->stream);
        n.commDestroy(ctx->ncclComm);
        ctx->ncclComm = nullptr;
    }
    ctx->rank = 0;
    ctx->size = 1;
    return VKRT_SUCCESS;
}

// Scene ingest for N ranks: the root has decoded and uploaded the images once (vkrt_textures_set); every
// other rank receives the device-resident texel pool over NVLink instead of decoding the same 69 files again.
//   1. a fixed-size header (count, then width/height per image) so that the receivers can lay out their pool;
//   2. the pool itself (all images, each 256-byte aligned - the layout vkrt_textures_set uses), one ncclBroadcast.
int broadcastTextures(Context* ctx, int root)
{
    constexpr uint32_t kMaxImages = 4095;
    if (ctx->size == 1)
    {
        return ctx->textureCount ? VKRT_SUCCESS : ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_textures_broadcast: no textures set");
    }
    if (!ctx->ncclComm)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_textures_broadcast: no communicator (vkrt_comm_init)");
    }
    NcclFns& n = nccl();
    cudaStream_t s = ctx->stream;
    const bool isRoot = ctx->rank == root;
    if (isRoot && (ctx->textureCount == 0 || ctx->textureCount > kMaxImages))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_textures_broadcast: the root needs 1..4095 textures set");
    }
    std::vector<uint32_t> header(2 * (kMaxImages + 1), 0u);
    if (isRoot)
    {
        header[0] = ctx->textureCount;
        for (uint32_t i = 0; i < ctx->textureCount; ++i)
        {
            header[2 + 2 * i] = ctx->textureHost[i].width;
            header[3 + 2 * i] = ctx->textureHost[i].height;
        }
    }
    DeviceBuffer hb;
    VKRT_CUDA_TRY(ctx, hb.reserve(header.size() * 4));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(hb.ptr, header.data(), header.size() * 4, cudaMemcpyHostToDevice, s));
    int rc = n.broadcast(hb.ptr, hb.ptr, header.size() * 4, kNcclUint8, root, ctx->ncclComm, s);
    if (rc == 0)
    {
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(header.data(), hb.ptr, header.size() * 4, cudaMemcpyDeviceToHost, s));
        VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    }
    hb.release();
    if (rc != 0)
    {
        return ctx->fail(VKRT_ERROR_NCCL, std::string("ncclBroadcast: ") + n.getErrorString(rc));
    }
    const uint32_t count = header[0];
    if (count == 0 || count > kMaxImages)
    {
        return ctx->fail(VKRT_ERROR_NCCL, "vkrt_textures_broadcast: bad header from the root");
    }
    std::vector<size_t> offsets(count);
    size_t total = 0;
    for (uint32_t i = 0; i < count; ++i)
    {
        offsets[i] = total;
        total += (((size_t)header[2 + 2 * i] * header[3 + 2 * i] * 4) + 255) & ~(size_t)255;
    }
    void* pool = nullptr;
    if (isRoot)
    {
        pool = ctx->textureAllocs.empty() ? nullptr : ctx->textureAllocs[0];
    }
    else
    {
        for (void* p : ctx->textureAllocs)
        {
            cudaFree(p);
        }
        ctx->textureAllocs.clear();
        ctx->textureCount = 0;
        VKRT_CUDA_TRY(ctx, cudaMalloc(&pool, total));
        ctx->textureAllocs.push_back(pool);
    }
    rc = n.broadcast(pool, pool, total, kNcclUint8, root, ctx->ncclComm, s);
    if (rc != 0)
    {
        return ctx->fail(VKRT_ERROR_NCCL, std::string("ncclBroadcast: ") + n.getErrorString(rc));
    }
    if (!isRoot)
    {
        std::vector<TextureRecord> recs(count);
        for (uint32_t i = 0; i < count; ++i)
        {
            recs[i].texels = static_cast<const uint8_t*>(pool) + offsets[i];
            recs[i].width = header[2 + 2 * i];
            recs[i].height = header[3 + 2 * i];
        }
        VKRT_CUDA_TRY(ctx, ctx->textureRecords.reserve(recs.size() * sizeof(TextureRecord)));
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->textureRecords.ptr, recs.data(), recs.size() * sizeof(TextureRecord), cudaMemcpyHostToDevice, s));
        VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
        ctx->textureCount = count;
        ctx->textureHost = recs;
        ctx->mipsHost.clear();
        ctx->mipsReady = false;
        ctx->materialsDirty = true;
        ctx->textureGeneration += 1;
    }
    else
    {
        VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    }
    return VKRT_SUCCESS;
}

int gatherTiles(Context* ctx, Frame* frame, int root)
{
    if (frame->size == 1)
    {
        frame->gathered = true;
        return VKRT_SUCCESS;
    }
    if (!ctx->ncclComm)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_tiles_gather: no communicator (vkrt_comm_init)");
    }
    if (frame->size != ctx->size || frame->rank != ctx->rank)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tiles_gather: frame was created under a different communicator");
    }
    NcclFns& n = nccl();
    // The exchange runs on its own stream, ordered after everything queued on the context stream so far.  The caller may
    // queue the next vkrt_trace at once: only its last kernel (k_resolve_sample, which rewrites the tile buffer and the
    // image) waits for this gather (joinGather), so the NCCL transfer and the untile overlap the next frame's traversal.
    if (!ctx->commStream)
    {
        VKRT_CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->commStream, cudaStreamNonBlocking));
        VKRT_CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->commReady, cudaEventDisableTiming));
    }
    if (!frame->gatherDone)
    {
        VKRT_CUDA_TRY(ctx, cudaEventCreateWithFlags(&frame->gatherDone, cudaEventDisableTiming));
    }
    if (int r = joinGather(ctx, frame)) // a second gather of the same frame without a trace in between
    {
        return r;
    }
    VKRT_CUDA_TRY(ctx, cudaEventRecord(ctx->commReady, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->commStream, ctx->commReady, 0));
    cudaStream_t s = ctx->commStream;
    const uint32_t tiles = frame->tilesX * frame->tilesY;
    const size_t tileBytes = (size_t)kTilePixels * 4;
    if (ctx->rank == root)
    {
        // staging for every peer's tiles, laid out rank after rank
        std::vector<size_t> offset((size_t)ctx->size + 1, 0);
        for (int r = 0; r < ctx->size; ++r)
        {
            offset[(size_t)r + 1] = offset[(size_t)r] + (r == root ? 0 : ownedTilesOf(tiles, r, ctx->size) * tileBytes);
        }
        VKRT_CUDA_TRY(ctx, ctx->gatherStaging.reserve(std::max<size_t>(offset[(size_t)ctx->size], 16)));
        uint8_t* staging = ctx->gatherStaging.as<uint8_t>();
        int rc = n.groupStart();
        for (int r = 0; r < ctx->size && rc == 0; ++r)
        {
            const size_t bytes = offset[(size_t)r + 1] - offset[(size_t)r];
            if (r != root && bytes)
            {
                rc = n.recv(staging + offset[(size_t)r], bytes, kNcclUint8, r, ctx->ncclComm, s);
            }
        }
        const int rc2 = n.groupEnd();
        if (rc != 0 || rc2 != 0)
        {
            return ctx->fail(VKRT_ERROR_NCCL, std::string("ncclRecv: ") + n.getErrorString(rc ? rc : rc2));
        }
        // one launch scatters the tiles of every peer: staged tile k of the concatenated peers maps back to its rank and
        // its index among that rank's tiles (ranks before the root hold one tile more when the tile count does not divide)
        {
            uint32_t stagedTiles = 0;
            for (int r = 0; r < ctx->size; ++r)
            {
                stagedTiles += r == root ? 0u : ownedTilesOf(tiles, r, ctx->size);
            }
            if (stagedTiles)
            {
                const uint32_t total = stagedTiles * (uint32_t)kTilePixels;
                const uint32_t grid = std::min<uint32_t>((total + 255) / 256, (uint32_t)ctx->smCount * 8u);
                k_untile_all<<<grid, 256, 0, s>>>(reinterpret_cast<const uint32_t*>(staging), stagedTiles, tiles, (uint32_t)root,
                                                 (uint32_t)ctx->size, frame->tilesX, frame->width, frame->height, frame->bgra8.as<uint32_t>());
                ctx->launches += 1;
            }
        }
        VKRT_CUDA_TRY(ctx, cudaGetLastError());
        frame->gathered = true;
    }
    else
    {
        const size_t bytes = (size_t)frame->ownedTiles * tileBytes;
        int rc = n.groupStart();
        if (rc == 0 && bytes)
        {
            rc = n.send(frame->tileBgra8.ptr, bytes, kNcclUint8, root, ctx->ncclComm, s);
        }
        const int rc2 = n.groupEnd();
        if (rc != 0 || rc2 != 0)
        {
            return ctx->fail(VKRT_ERROR_NCCL, std::string("ncclSend: ") + n.getErrorString(rc ? rc : rc2));
        }
    }
    VKRT_CUDA_TRY(ctx, cudaEventRecord(frame->gatherDone, ctx->commStream));
    frame->gatherPending = true;
    return VKRT_SUCCESS;
}

int joinGather(Context* ctx, Frame* frame)
{
    if (frame->gatherPending)
    {
        VKRT_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, frame->gatherDone, 0));
        frame->gatherPending = false;
    }
    return VKRT_SUCCESS;
}

} // namespace vkrt
