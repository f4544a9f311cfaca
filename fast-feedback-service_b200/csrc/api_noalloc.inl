// api_noalloc.inl -- ONE entry point that allocates nothing (no counterpart in the reference).
// Some callers cannot call cudaMalloc: code under CUDA
// graph capture, tight serving loops, harnesses that forbid it.  ffs_get_strong_u8 (api_core.inl)
// unpacks a slot's 1-bit strong plane into a 0/1 byte plane through a cudaMalloc'ed scratch buffer;
// ffs_get_strong_u8_static does the same through a file-scope device array of fixed size.
//
//   * g_unpack_scratch lives in every GPU that loads the library until the process exits
//     (FFS_NOALLOC_MAX_PIXELS bytes = 4.5 MiB: Eiger2 4M 2070x2167 and Jungfrau 4M 2068x2164 fit,
//     Eiger2 16M does not and is rejected with FFS_ERR_CAPACITY).
//   * In host code the array's NAME is a host-side stand-in, not a device address: its address is
//     taken once per device with cudaGetSymbolAddress, that device current, and the pointer is
//     what the kernel gets.
//   * Every call on a device shares the array and it keeps the last call's contents.  The unpack
//     kernel writes every byte of [0, width*height) before anything reads it, so nothing has to
//     start at zero; calls that overlap on one GPU (two host threads, two contexts on one device)
//     race on it -- the entry point serialises its own callers with a mutex, and waits for its
//     stream before it returns, but it cannot see other processes' streams.
static __device__ __align__(256) uint8_t g_unpack_scratch[FFS_NOALLOC_MAX_PIXELS];

static std::mutex g_noalloc_mu;
static uint8_t* g_unpack_scratch_dev[64] = {};  // per device ordinal: device address of the array

extern "C" size_t ffs_noalloc_max_pixels(void) { return (size_t)FFS_NOALLOC_MAX_PIXELS; }

extern "C" int ffs_get_strong_u8_static(ffs_ctx* c, int slot, uint8_t* out) {
    if (!slots_ok(c, slot, 1) || !out) return FFS_ERR_ARG;
    const size_t px = (size_t)c->p.width * c->p.height;
    if (px > (size_t)FFS_NOALLOC_MAX_PIXELS) {
        c->err = "ffs_get_strong_u8_static: " + std::to_string(px) + " pixels exceed the static scratch of " +
                 std::to_string((size_t)FFS_NOALLOC_MAX_PIXELS) + " (FFS_NOALLOC_MAX_PIXELS); use ffs_get_strong_u8";
        return FFS_ERR_CAPACITY;
    }
    if (c->p.device < 0 || c->p.device >= 64) return FFS_ERR_ARG;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);  // callers of this entry point share one array per GPU
    FFS_CK(c, cudaSetDevice(c->p.device));
    uint8_t*& scratch = g_unpack_scratch_dev[c->p.device];
    if (!scratch) {  // set-up, once per device, with that device current
        void* ptr = nullptr;
        FFS_CK(c, cudaGetSymbolAddress(&ptr, g_unpack_scratch));
        scratch = static_cast<uint8_t*>(ptr);
    }
    FFS_CK(c, cudaDeviceSynchronize());  // the slot's threshold may still be running on another stream
    unpack_bits_kernel<<<(unsigned)((px + 255) / 256), 256, 0, c->stream>>>(
      reinterpret_cast<const uint32_t*>(c->d_strong + (size_t)slot * c->zstride), c->bits_pitch, (int)c->p.width,
      (int)c->p.height, scratch);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    FFS_CK(c, cudaMemcpyAsync(out, scratch, px, cudaMemcpyDeviceToHost, c->stream));
    FFS_CK(c, cudaStreamSynchronize(c->stream));
    return FFS_OK;
}
