// The product engine: the C ABI of libfsq_b200.so (include/fsq.h).  No CPU fallback: without the library or a device
// the run fails with the library's message.
#include "engine.hh"
#include "../include/fsq.h"

namespace fsq_host {

char const *engine_name() { return fsq_version(); }

engine_status run_engine(std::uint8_t const *const *rows, std::uint32_t m, std::uint64_t n, std::uint8_t const comp[256], std::uint64_t L, segmentation &out, std::string &message)
{
	fsq_ctx *ctx = nullptr;
	if (fsq_create(0, &ctx)) { message = fsq_last_error(nullptr); return ENGINE_ERROR; }
	// ctx, msa and seg stay alive until everything that reads them (here: the copies below) is done
	struct cleanup
	{
		fsq_ctx *ctx; fsq_msa *msa{}; fsq_segmentation *seg{};
		~cleanup() { fsq_segmentation_free(seg); fsq_msa_free(msa); fsq_destroy(ctx); }
	} c{ctx};
	auto failed = [&] { message = fsq_last_error(ctx); return ENGINE_ERROR; };
	if (fsq_msa_from_rows(ctx, rows, m, n, comp, &c.msa)) return failed();
	int rc = fsq_segment(ctx, c.msa, L, &c.seg);
	if (FSQ_ERR_SHORT == rc)
	{
		// generate_context.cc:379-390 -> segmentation_sp_context (n < 2L)
		out.short_path = true;
		out.row_of_run.resize(m); out.run_len.resize(m);
		std::uint32_t count = 0;
		if (fsq_distinct_rows(ctx, c.msa, out.row_of_run.data(), out.run_len.data(), &count)) return failed();
		out.row_of_run.resize(count); out.run_len.resize(count);
		out.K = count;
		return ENGINE_OK;
	}
	if (FSQ_ERR_NOT_REDUCIBLE == rc) { message = fsq_last_error(ctx); return ENGINE_NOT_REDUCIBLE; }
	if (rc) return failed();
	std::uint32_t const count = fsq_segmentation_count(c.seg);
	out.K = fsq_segmentation_max_segment_size(c.seg);
	out.lb.resize(count); out.rb.resize(count); out.size.resize(count);
	out.a.resize((std::size_t) count * m); out.d.resize((std::size_t) count * m);
	if (fsq_segmentation_bounds(c.seg, out.lb.data(), out.rb.data(), out.size.data())) return failed();
	if (fsq_segmentation_states(c.seg, out.a.data(), out.d.data())) return failed();
	return ENGINE_OK;
}

} // namespace fsq_host
