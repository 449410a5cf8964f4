// The reference writes this include with Windows path separators (src/IO_flow.cpp:4); on Linux the
// name is looked up literally, so this file carries that literal name and forwards to the real header.
#include <IO_flow.hpp>
