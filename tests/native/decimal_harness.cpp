// Host build of csrc/decimal.cuh for the CPU unit test (tests/test_decimal.py).
#include "../../pyspark-tf-gke_b200/csrc/decimal.cuh"
extern "C" __attribute__((visibility("default"))) int dec_parse_host(const unsigned char* p, int len, double* out,
                                                                      int* flags) {
    return etl::dec_parse(p, len, out, flags);
}
// parse `count` NUL-free strings packed back to back (offsets[count+1]); returns statuses/values
extern "C" __attribute__((visibility("default"))) void dec_parse_many(const unsigned char* buf, const long long* offsets,
                                                                       long long count, double* out, int* status,
                                                                       int* flags) {
    for (long long i = 0; i < count; ++i)
        status[i] = etl::dec_parse(buf + offsets[i], (int)(offsets[i + 1] - offsets[i]), out + i, flags + i);
}
