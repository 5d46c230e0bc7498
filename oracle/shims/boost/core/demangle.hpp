// Test-infrastructure shim (oracle/_ref build only). Not product code.
#pragma once
#include <string>
namespace boost { namespace core { inline std::string demangle(char const *n) { return n; } } }
