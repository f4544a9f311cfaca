static int stream_setup(ffs_ctx*) { return FFS_OK; }
static void stream_teardown(ffs_ctx*) {}
extern "C" int ffs_acquire(ffs_ctx*, int*, void**) { return FFS_ERR_STATE; }
extern "C" int ffs_submit(ffs_ctx*, int, uint64_t) { return FFS_ERR_STATE; }
extern "C" int ffs_submit_copy(ffs_ctx*, uint64_t, const void*) { return FFS_ERR_STATE; }
extern "C" int ffs_wait(ffs_ctx*, ffs_frame_result*, ffs_spot*, size_t, size_t*) { return FFS_ERR_STATE; }
extern "C" int ffs_pending(ffs_ctx*) { return 0; }
