static int extended_setup(ffs_ctx*) { return FFS_OK; }
static int extended_impl(ffs_ctx* c, int, int, cudaStream_t) { c->err = "extended: not built yet"; return FFS_ERR_STATE; }
static int cc_setup(ffs_ctx*) { return FFS_OK; }
static void cc_teardown(ffs_ctx*) {}
extern "C" int ffs_process(ffs_ctx* c, int, int, void*) { return FFS_ERR_STATE; }
extern "C" int ffs_fetch(ffs_ctx*, int, ffs_frame_result*, ffs_spot*, size_t, size_t*) { return FFS_ERR_STATE; }
extern "C" int ffs_get_labels(ffs_ctx*, int, int64_t*) { return FFS_ERR_STATE; }
extern "C" int ffs_sweep_begin(ffs_ctx*, uint64_t) { return FFS_ERR_STATE; }
extern "C" int ffs_sweep_finish(ffs_ctx*, ffs_spot**, size_t*) { return FFS_ERR_STATE; }
extern "C" int ffs_sweep_export_plane(ffs_ctx*, int, uint32_t**, size_t*) { return FFS_ERR_STATE; }
extern "C" size_t ffs_filter_spots(ffs_spot*, size_t, uint32_t, float, size_t*, size_t*) { return 0; }
