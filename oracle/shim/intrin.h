// oracle/shim/intrin.h — TEST INFRASTRUCTURE. Empty stand-in for the MSVC-only header that the
// reference includes (main.cpp:5) but never uses.
