// bz_error.h -- per-thread error text behind bz_last_error() (defined in csrc/bz_capi.cu)
#ifndef BZ_ERROR_H
#define BZ_ERROR_H

#include <string>

int bz_fail(int code, const std::string &text);

#endif
