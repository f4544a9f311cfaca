static int extended_setup(ffs_ctx*) { return FFS_OK; }
static int extended_impl(ffs_ctx* c, int, int, cudaStream_t) { c->err = "extended: not built yet"; return FFS_ERR_STATE; }
