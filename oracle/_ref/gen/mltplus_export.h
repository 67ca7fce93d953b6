#define MLTPLUS_EXPORT __attribute__((visibility("default")))
