// TEST INFRASTRUCTURE: the four build-info constants the reference's tools expect (common/common.h:24-27); the reference
// generates them with cmake from common/build-info.cpp.in, which oracle/Makefile.ref does not run.
int          LLAMA_BUILD_NUMBER = 0;
char const * LLAMA_COMMIT       = "9561c161";
char const * LLAMA_COMPILER     = "g++ (oracle/Makefile.ref)";
char const * LLAMA_BUILD_TARGET = "x86_64-linux-gnu";
