// Exception.h -- error convention of the reference host code (src/Exception.h:31-80): failures throw std::runtime_error.
#pragma once
#include <cstdio>
#include <stdexcept>
#include <string>

#include "grt.h"

#define GRT_CHECK(ctx, call)                                                                              \
    do {                                                                                                  \
        if ((call) != 0) {                                                                                \
            char grt_msg_[1024];                                                                          \
            snprintf(grt_msg_, sizeof(grt_msg_), "libgrt call '%s' failed: %s (%s:%d)", #call, grtGetLastError(ctx), __FILE__, __LINE__); \
            throw std::runtime_error(grt_msg_);                                                           \
        }                                                                                                 \
    } while (0)
