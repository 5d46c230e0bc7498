// Test-infrastructure shim (oracle/_ref build only). Not product code.
#pragma once
#include <algorithm>
namespace boost { namespace sort {
	template <typename I> void block_indirect_sort(I b, I e) { std::sort(b, e); }
	template <typename I, typename C> void block_indirect_sort(I b, I e, C c) { std::sort(b, e, c); }
} }
