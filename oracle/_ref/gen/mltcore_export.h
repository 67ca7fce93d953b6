#define MLTCORE_EXPORT __attribute__((visibility("default")))
