/* Declaration-only stand-in for Barsdell's dedisp library header.
 *
 * TEST INFRASTRUCTURE ONLY.  The reference's utils/exceptions.hpp:8 includes
 * "dedisp.h", so every reference translation unit needs one on the include
 * path even when (as here) no dedispersion is performed.  This file declares
 * just the error type that header uses; it contains no dedisp code.
 */
#pragma once
typedef int dedisp_error;
enum { DEDISP_NO_ERROR = 0 };
static inline const char* dedisp_get_error_string(dedisp_error) { return "dedisp not linked (oracle shim)"; }
