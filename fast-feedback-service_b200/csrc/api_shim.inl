extern "C" void* spotfinder_create(size_t, size_t) { return nullptr; }
extern "C" void spotfinder_free(void*) {}
extern "C" uint32_t spotfinder_standard_dispersion(void*, ffs_image_t*, ffs_bool**) { return 0; }
extern "C" void* spotfinder_create32(size_t, size_t) { return nullptr; }
extern "C" uint32_t spotfinder_standard_dispersion32(void*, ffs_image_t*, ffs_bool**) { return 0; }
