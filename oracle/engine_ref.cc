// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  The host driver's engine seam (host/engine.hh) filled by the reference's own
// arithmetic compiled in place (ref_harness.cc: pbwt_context, SDSL RMQ, rmq.hh + the restated block-based driver).  The
// binary built from it (oracle/_ref/founder_sequences_ref) writes the founders / segments files that the product's
// (host/founder_sequences_b200, CUDA through the C ABI) are compared with byte for byte (tests/test_host_driver.py).
#include "../host/engine.hh"
#include <cstdint>
#include <vector>

extern "C" {
void *fsqref_create(std::uint8_t const *rows, std::uint32_t m, std::uint64_t n);
void fsqref_destroy(void *h);
int fsqref_segment(void *h, std::uint64_t L, double sample_rate_multiplier, int with_texts);
std::uint32_t fsqref_K(void *h);
std::uint64_t fsqref_reduced_size(void *h);
void fsqref_get_reduced(void *h, std::uint64_t *lb, std::uint64_t *rb, std::uint32_t *mx, std::uint32_t *sz);
void fsqref_get_reduced_states(void *h, std::uint32_t *a, std::uint32_t *d);
std::uint64_t fsqref_short_path(void *h, std::uint32_t *row, std::uint32_t *len);
}

namespace fsq_host {

char const *engine_name() { return "reference arithmetic (oracle/_ref)"; }

engine_status run_engine(std::uint8_t const *const *rows, std::uint32_t m, std::uint64_t n, std::uint8_t const comp[256], std::uint64_t L, segmentation &out, std::string &message)
{
	std::vector <std::uint8_t> flat((std::size_t) m * n);
	for (std::uint32_t r = 0; r < m; ++r) std::copy(rows[r], rows[r] + n, flat.begin() + (std::size_t) r * n);
	void *h = fsqref_create(flat.data(), m, n);
	struct cleanup { void *h; ~cleanup() { fsqref_destroy(h); } } c{h};
	int const status = fsqref_segment(h, L, 4.0, 0);
	if (1 == status)
	{
		out.short_path = true;
		out.row_of_run.resize(m); out.run_len.resize(m);
		std::uint64_t const count = fsqref_short_path(h, out.row_of_run.data(), out.run_len.data());
		out.row_of_run.resize(count); out.run_len.resize(count);
		out.K = (std::uint32_t) count;
		return ENGINE_OK;
	}
	if (2 == status) { message = "the maximum segment size is equal to the number of input sequences"; return ENGINE_NOT_REDUCIBLE; }
	if (status) { message = "reference harness failed"; return ENGINE_ERROR; }
	std::uint64_t const count = fsqref_reduced_size(h);
	out.K = fsqref_K(h);
	out.lb.resize(count); out.rb.resize(count); out.size.resize(count);
	std::vector <std::uint32_t> mx(count);
	out.a.resize((std::size_t) count * m); out.d.resize((std::size_t) count * m);
	fsqref_get_reduced(h, out.lb.data(), out.rb.data(), mx.data(), out.size.data());
	fsqref_get_reduced_states(h, out.a.data(), out.d.data());
	return ENGINE_OK;
}

} // namespace fsq_host
