// api_shim.inl -- drop-in for the reference's CPU C API (baseline/spotfinder/baseline.h:9-13,
// implementation baseline.cpp:849-881): same names, same argument meaning, same ownership
// (the returned bool plane is owned by the context and valid until the next call).  The
// reference passes the mask with every image; we keep it resident and re-upload on change.

namespace {
struct ShimCtx {
    size_t width, height;
    uint32_t bytes_per_pixel;
    ffs_ctx* ctx = nullptr;
    std::vector<uint8_t> mask;   // last mask seen
    std::vector<uint8_t> result; // bool plane handed to the caller
};

void* shim_create(size_t width, size_t height, uint32_t bpp) {
    ShimCtx* s = new (std::nothrow) ShimCtx();
    if (!s) return nullptr;
    s->width = width;
    s->height = height;
    s->bytes_per_pixel = bpp;
    ffs_params p;
    ffs_default_params(&p);
    p.width = (uint32_t)width;
    p.height = (uint32_t)height;
    p.bytes_per_pixel = bpp;
    p.n_slots = 1;
    if (ffs_create(&p, nullptr, &s->ctx) != FFS_OK) {
        fprintf(stderr, "spotfinder_create: %s\n", ffs_last_error(nullptr));
        delete s;
        return nullptr;
    }
    s->result.resize(width * height);
    return s;
}

uint32_t shim_dispersion(void* context, ffs_image_t* image, ffs_bool** destination) {
    ShimCtx* s = static_cast<ShimCtx*>(context);
    if (!s || !image || !image->data || image->fast != s->width || image->slow != s->height) return 0;
    ffs_ctx* c = s->ctx;
    const size_t px = s->width * s->height;
    if (image->mask && (s->mask.size() != px || memcmp(s->mask.data(), image->mask, px) != 0)) {
        s->mask.assign(image->mask, image->mask + px);
        if (cudaMemcpy(c->d_mask, image->mask, px, cudaMemcpyHostToDevice) != cudaSuccess) return 0;
        if (repack_mask(c) != FFS_OK) return 0;
    }
    if (ffs_upload(c, 0, 1, image->data) != FFS_OK) return 0;
    if (ffs_threshold(c, 0, 1, nullptr) != FFS_OK) return 0;
    if (ffs_get_strong_u8(c, 0, s->result.data()) != FFS_OK) return 0;
    unsigned count = 0;
    cudaMemcpy(&count, c->d_strong_count, sizeof(count), cudaMemcpyDeviceToHost);
    static_assert(sizeof(ffs_bool) == 1, "bool plane is one byte per pixel");
    if (destination) *destination = reinterpret_cast<ffs_bool*>(s->result.data());
    return count;
}
}  // namespace

extern "C" void* spotfinder_create(size_t width, size_t height) { return shim_create(width, height, 2); }
extern "C" void* spotfinder_create32(size_t width, size_t height) { return shim_create(width, height, 4); }
extern "C" void spotfinder_free(void* context) {
    ShimCtx* s = static_cast<ShimCtx*>(context);
    if (!s) return;
    ffs_destroy(s->ctx);
    delete s;
}
extern "C" uint32_t spotfinder_standard_dispersion(void* context, ffs_image_t* image, ffs_bool** destination) {
    return shim_dispersion(context, image, destination);
}
extern "C" uint32_t spotfinder_standard_dispersion32(void* context, ffs_image_t* image, ffs_bool** destination) {
    return shim_dispersion(context, image, destination);
}

// Host-side ingestion helper shared with the spotfinder executable (cli/bslz4.hpp): decode one
// bitshuffle-LZ4 HDF5-filter chunk (12-byte header + body) into n dense elements.
#include "cli/bslz4.hpp"
extern "C" int ffs_bslz4_decode_chunk(const void* chunk, size_t chunk_bytes, void* out, size_t n_elements,
                                      size_t elem_size) {
    if (!chunk || !out || (elem_size != 2 && elem_size != 4)) return FFS_ERR_ARG;
    static thread_local std::vector<uint8_t> scratch;
    return bslz4::decompress_chunk(static_cast<const uint8_t*>(chunk), chunk_bytes, out, n_elements, elem_size,
                                   scratch) == 0
             ? FFS_OK
             : FFS_ERR_ARG;
}
