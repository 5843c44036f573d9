extern "C" __attribute__((visibility("default"))) int dmp_rt_version() { return 100; }
