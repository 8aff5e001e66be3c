// Declaration-only stand-in so /root/reference/src/ai/vision.cc compiles when the
// un-vendored stb header (MODULE.bazel:38) is absent. Only the libtorch overloads of
// vision.cc are used by the oracle; the stb wrapper (vision.cc:86-95) is NOT an oracle.
#pragma once
typedef enum { STBIR_1CHANNEL = 1 } stbir_pixel_layout;
extern "C" unsigned char *stbir_resize_uint8_linear(const unsigned char *, int, int, int,
                                                    unsigned char *, int, int, int,
                                                    stbir_pixel_layout);
