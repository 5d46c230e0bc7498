// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// oracle/_ref/libfsq_ref.so: the reference's own arithmetic, compiled unmodified
// from the sources where they lie under /root/reference (never copied), behind a
// small C ABI so that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
// leg can use it as the checker / CPU baseline.  Only those three may load it.
//
// Compiled from the reference as-is (through #include / the Makefile's source list):
//   libbio::pbwt::pbwt_context  + pbwt.hh free functions + array_list   (lib/libbio/include/libbio/pbwt_context.hh, pbwt.hh, array_list.hh)
//   sdsl::range_maximum_sct     (lib/sdsl-lite)
//   founder_sequences::rmq      (include/founder_sequences/rmq.hh)
//   founder_sequences::segmentation_dp_arg (include/founder_sequences/segmentation_dp_arg.hh)
//   founder_sequences::create_segment_texts_task::execute (founder-sequences/create_segment_texts_task.cc)
//   libbio::consecutive_alphabet_as_builder (lib/libbio/include/libbio/consecutive_alphabet.hh)
//   libbio::set_intersection_size (lib/libbio/include/libbio/algorithm/set_count.hh)
//
// Restated here (the originals need Clang Blocks / libdispatch / LEMON, which this image lacks;
// SURVEY.md F7), each following the cited lines and *calling* the compiled pieces above:
//   first pass staging            founder-sequences/segmentation_lp_context.cc:26-188
//   DP inner function             founder-sequences/segmentation_lp_context.cc:393-481
//   traceback                     founder-sequences/segmentation_lp_context.cc:191-224
//   sample re-scan                founder-sequences/segmentation_lp_context.cc:229-332, update_pbwt_task.cc:13-35
//   greedy reduction              founder-sequences/segmentation_lp_context.cc:335-390
//   run extraction (bipartite)    founder-sequences/join_context.cc:63-126
//   edge weights                  founder-sequences/merge_segments_task.cc:133-176

#include <founder_sequences/founder_sequences.hh>
#include <founder_sequences/rmq.hh>
#include <founder_sequences/segmentation_dp_arg.hh>
#include <founder_sequences/segment_text.hh>
#include <founder_sequences/create_segment_texts_task.hh>
#include <libbio/algorithm.hh>
#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>

namespace fseq = founder_sequences;
namespace lb = libbio;

// libbio::dispatch_ptr pulls these two symbols in; nothing from libdispatch is used.
#ifndef FSQ_REF_HARNESS_NO_DISPATCH_STUBS      // (oracle/_ref/founder_sequences_ref gets them from the host driver's source)
extern "C" void dispatch_retain(dispatch_object_t) {}
extern "C" void dispatch_release(dispatch_object_t) {}
#endif

namespace {

	// pbwt_context's members are protected with no setters; see SURVEY.md §8b.
	struct open_sample : public fseq::pbwt_context
	{
		using fseq::pbwt_context::pbwt_context;
		open_sample(fseq::pbwt_context const &other): fseq::pbwt_context(other) {}
		std::size_t &seq_idx() { return m_sequence_idx; }
	};

	struct harness
	{
		std::uint32_t							m{};
		std::uint64_t							n{};
		std::vector <std::uint8_t>				bytes;		// Row-major copy of the input.
		fseq::sequence_vector					sequences;
		fseq::alphabet_type						alphabet;

		// Results of segment().
		std::uint64_t							L{};
		fseq::segmentation_traceback_vector	dp;			// n - L + 1 entries
		fseq::segmentation_traceback_vector	traceback;
		std::uint32_t							K{};
		std::vector <fseq::pbwt_sample_type>	traceback_states;	// (a_rb, d_rb) per traceback entry
		fseq::segmentation_traceback_vector	reduced;
		std::vector <fseq::pbwt_sample_type>	reduced_states;
		fseq::segment_text_matrix				segment_texts;		// bipartite: K slots per reduced segment
		double									first_pass_seconds{};
	};

	// Restatement of calculate_segmentation_lp_dp_arg, segmentation_lp_context.cc:393-481.
	void dp_arg(
		typename fseq::pbwt_context::divergence_count_list const &counts,
		fseq::segmentation_traceback_vector const &tb,
		fseq::segmentation_traceback_vector_rmq const &tb_rmq,
		std::size_t const seq_count,
		std::size_t const L,
		std::size_t const lb_,
		std::size_t const text_pos,
		fseq::segmentation_dp_arg &min_arg
	)
	{
		auto it(counts.cbegin_pairs());
		auto const end(counts.cend_pairs());
		std::size_t diff(it->second);
		std::size_t dp_lb(lb_);
		std::size_t dp_rb(it->first);
		if (lb_ == dp_rb)
		{
			auto const sz(seq_count - diff);
			fseq::segmentation_dp_arg const cur(lb_, 1 + text_pos, sz, sz);
			if (cur < min_arg)
				min_arg = cur;
			++it;
			dp_rb = it->first;
			diff += it->second;
		}
		++it;
		while (it != end)
		{
			dp_lb = dp_rb;
			dp_rb = it->first;
			std::size_t dp_rb_c(dp_rb);
			bool ok(true);
			if (text_pos + 2 - L < dp_rb_c)
				dp_rb_c = text_pos + 2 - L;
			if (dp_lb < lb_ + L)
			{
				if (lb_ + L < dp_rb_c)
					dp_lb = lb_ + L;
				else
					ok = false;
			}
			if (ok && dp_lb < dp_rb_c)
			{
				auto const idx(tb_rmq(dp_lb - L, dp_rb_c - L));
				auto const lhs(tb[idx].segment_max_size);
				auto const rhs(seq_count - diff);
				fseq::segmentation_dp_arg cur(idx + L, 1 + text_pos, lb::max_ct(lhs, rhs), rhs);
				if (cur < min_arg)
					min_arg = cur;
			}
			diff += it->second;
			++it;
		}
	}
}


extern "C" {

	void *fsqref_create(std::uint8_t const *rows, std::uint32_t m, std::uint64_t n)
	{
		auto *h(new harness);
		h->m = m;
		h->n = n;
		h->bytes.assign(rows, rows + std::size_t(m) * n);
		for (std::uint32_t i(0); i < m; ++i)
			h->sequences.emplace_back(h->bytes.data() + std::size_t(i) * n, n);

		// generate_context.cc:134-147
		lb::consecutive_alphabet_as_builder <std::uint8_t> builder;
		builder.init();
		for (auto const &vec : h->sequences)
			builder.prepare(vec);
		builder.compress();
		using std::swap;
		swap(h->alphabet, builder.alphabet());
		return h;
	}

	void fsqref_destroy(void *hp) { delete static_cast <harness *>(hp); }

	// comp[c] for every byte value, as the reference's alphabet maps it.
	void fsqref_alphabet(void *hp, std::uint8_t *comp256)
	{
		auto *h(static_cast <harness *>(hp));
		for (unsigned c(0); c < 256; ++c)
			comp256[c] = h->alphabet.char_to_comp(c);
	}

	// Run the reference PBWT over columns [0, ncols) and dump a_{k+1}, d_{k+1} after every column k
	// (ncols × m each) and the histogram pairs.  hist_off has ncols+1 entries; hist_v / hist_c are
	// caller-allocated with capacity hist_cap pairs.  Returns the number of pairs written, or -1 on overflow.
	long long fsqref_scan_dump(
		void *hp, std::uint64_t ncols,
		std::uint32_t *a_out, std::uint32_t *d_out,
		std::uint64_t *hist_off, std::uint32_t *hist_v, std::uint32_t *hist_c, std::uint64_t hist_cap
	)
	{
		auto *h(static_cast <harness *>(hp));
		fseq::pbwt_context ctx(h->sequences, h->alphabet);
		ctx.prepare();
		std::uint64_t pairs(0);
		bool overflow(false);
		if (hist_off) hist_off[0] = 0;
		ctx.process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(ncols, [&](){
			auto const k(ctx.sequence_idx());
			auto const &a(ctx.output_permutation());
			auto const &d(ctx.output_divergence());
			for (std::uint32_t i(0); i < h->m; ++i)
			{
				if (a_out) a_out[k * h->m + i] = a[i];
				if (d_out) d_out[k * h->m + i] = d[i];
			}
			if (hist_off)
			{
				auto const &counts(ctx.output_divergence_value_counts());
				for (auto it(counts.cbegin_pairs()), end(counts.cend_pairs()); it != end; ++it)
				{
					if (pairs < hist_cap)
					{
						hist_v[pairs] = it->first;
						hist_c[pairs] = it->second;
					}
					else
						overflow = true;
					++pairs;
				}
				hist_off[k + 1] = pairs;
			}
		});
		return overflow ? -1 : (long long) pairs;
	}

	// First pass only over the first ncols columns (PBWT + histogram + DP per column, sampling as the
	// reference does), timed.  Used as the "reference" CPU baseline.  Returns seconds.
	double fsqref_time_first_pass(void *hp, std::uint64_t L, std::uint64_t ncols, double sample_rate_multiplier)
	{
		auto *h(static_cast <harness *>(hp));
		auto const n(std::min <std::uint64_t>(ncols, h->n));
		auto const seq_count(h->m);
		fseq::segmentation_traceback_vector tb(n - L + 1);
		fseq::segmentation_traceback_vector_rmq tb_rmq(tb);
		fseq::pbwt_context ctx(h->sequences, h->alphabet, lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS);
		if (0 < sample_rate_multiplier)
			ctx.set_sample_rate(std::ceil(sample_rate_multiplier * std::sqrt(double(h->n))));
		auto const t0(std::chrono::steady_clock::now());
		ctx.prepare();
		ctx.process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(n - L, [&](){
			auto const idx(ctx.sequence_idx());
			if (idx + 1 < L)
				return;
			fseq::segmentation_dp_arg min_arg(0, 1 + idx, seq_count, seq_count);
			dp_arg(ctx.output_divergence_value_counts(), tb, tb_rmq, seq_count, L, 0, idx, min_arg);
			auto const tb_idx(idx + 1 - L);
			tb[tb_idx] = min_arg;
			tb_rmq.update(tb_idx);
		});
		auto const t1(std::chrono::steady_clock::now());
		return std::chrono::duration <double>(t1 - t0).count();
	}

	// The whole long path: first pass, traceback, re-scan, greedy reduction, segment texts.
	// Returns 0 on success, 1 if n < 2L (short path), 2 if K == m (the reference exits with failure).
	int fsqref_segment(void *hp, std::uint64_t L, double sample_rate_multiplier, int with_texts)
	{
		auto *h(static_cast <harness *>(hp));
		auto const n(h->n);
		auto const seq_count(h->m);
		if (n < 2 * L)
			return 1;

		h->L = L;
		h->dp.clear();
		h->traceback.clear();
		h->traceback_states.clear();
		h->reduced.clear();
		h->reduced_states.clear();
		h->segment_texts.clear();

		// generate_context.cc:122
		std::uint64_t const rate(0 < sample_rate_multiplier ? std::ceil(sample_rate_multiplier * std::sqrt(double(n))) : 1 + n);

		auto const t0(std::chrono::steady_clock::now());

		// segmentation_lp_context.cc:26-58
		fseq::pbwt_context ctx(h->sequences, h->alphabet, lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS);	// segmentation_lp_context.hh:102
		ctx.set_sample_rate(rate);
		ctx.prepare();
		auto const dp_size(n - L + 1);
		h->dp.resize(dp_size);
		auto &tb(h->dp);
		fseq::segmentation_traceback_vector_rmq tb_rmq(tb);
		ctx.process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(L - 1, [](){});

		// :65-98
		{
			auto const limit(std::min(2 * L, n - L) - 1);
			ctx.process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(limit, [&](){
				auto const idx(ctx.sequence_idx());
				auto const &counts(ctx.output_divergence_value_counts());
				std::size_t diff(0);
				auto const begin(counts.cbegin_pairs());
				if (counts.cend_pairs() != begin && 0 == begin->first)
					diff = begin->second;
				auto const tb_idx(idx + 1 - L);
				auto const sz(seq_count - diff);
				tb[tb_idx] = fseq::segmentation_dp_arg(0, 1 + idx, sz, sz);
				tb_rmq.update(tb_idx);
			});
		}

		// :106-142
		ctx.process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(n - L, [&](){
			auto const idx(ctx.sequence_idx());
			fseq::segmentation_dp_arg min_arg(0, 1 + idx, seq_count, seq_count);
			dp_arg(ctx.output_divergence_value_counts(), tb, tb_rmq, seq_count, L, 0, idx, min_arg);
			auto const tb_idx(idx + 1 - L);
			tb[tb_idx] = min_arg;
			tb_rmq.update(tb_idx);
		});

		// :149-185
		ctx.process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(n, [](){});
		{
			auto counts(ctx.last_divergence_value_counts());
			auto const idx(ctx.sequence_idx());
			fseq::segmentation_dp_arg min_arg(0, idx, seq_count, seq_count);
			dp_arg(counts, tb, tb_rmq, seq_count, L, 0, idx - 1, min_arg);
			tb[dp_size - 1] = min_arg;
		}

		h->first_pass_seconds = std::chrono::duration <double>(std::chrono::steady_clock::now() - t0).count();

		// :191-224
		{
			std::size_t arg_idx(dp_size - 1);
			while (true)
			{
				auto const &cur(tb[arg_idx]);
				h->traceback.push_back(cur);
				if (0 == cur.lb)
					break;
				arg_idx = cur.lb - L;
			}
			std::reverse(h->traceback.begin(), h->traceback.end());
			h->K = h->traceback.back().segment_max_size;
		}

		// generate_context.cc:192-200
		if (! (h->K < seq_count))
			return 2;

		// :229-332 and update_pbwt_task.cc:13-35.  For every traceback right bound, re-run the PBWT from
		// the last sample at or before it; (a_rb, d_rb) do not depend on which sample is used.
		{
			auto &samples(ctx.samples());
			std::size_t si(0);
			std::unique_ptr <open_sample> cur;
			std::size_t cur_si(SIZE_MAX);
			for (auto const &arg : h->traceback)
			{
				while (si + 1 < samples.size() && samples[si + 1].sequence_idx() <= arg.rb)
					++si;
				if (samples.empty() || arg.rb < samples[si].sequence_idx())
				{
					// Before the first sample (only possible for rb == 0, which cannot happen as L ≥ 1).
					return 3;
				}
				if (cur_si != si)
				{
					cur.reset(new open_sample(samples[si]));
					cur->set_sample_rate(std::numeric_limits <std::uint64_t>::max());
					cur_si = si;
				}
				cur->process <lb::pbwt::context_field::DIVERGENCE_VALUE_COUNTS>(arg.rb, [](){});
				auto &copied(h->traceback_states.emplace_back());
				copied.set_fields_in_use(static_cast <lb::pbwt::context_field>(
					lb::pbwt::context_field::INPUT_PERMUTATION | lb::pbwt::context_field::INPUT_DIVERGENCE));
				copied.copy_fields_in_use(*cur);
			}
		}

		// :335-390
		{
			std::size_t current_lb(0);
			std::size_t prev_size(h->traceback.front().segment_size);
			std::size_t prev(0);
			for (std::size_t t(1); t < h->traceback.size(); ++t)
			{
				auto const &sample(h->traceback_states[t]);
				auto const &arg(h->traceback[t]);
				auto const sample_size(sample.unique_substring_count_lhs(current_lb));
				if (sample_size <= h->K)
					prev_size = sample_size;
				else
				{
					h->reduced.emplace_back(current_lb, h->traceback_states[prev].sequence_idx(), prev_size);
					prev_size = arg.segment_size;
					current_lb = h->traceback_states[prev].sequence_idx();
					h->reduced_states.emplace_back(h->traceback_states[prev]);
				}
				prev = t;
			}
			h->reduced.emplace_back(current_lb, h->traceback_states[prev].sequence_idx(), prev_size);
			h->reduced_states.emplace_back(h->traceback_states[prev]);
		}

		// join_context.cc:63-126 (bipartite branch) + bipartite_matcher.cc:28-44 → create_segment_texts_task (compiled).
		if (with_texts)
		{
			h->segment_texts.resize(h->reduced.size());
			for (std::size_t s(0); s < h->reduced.size(); ++s)
			{
				fseq::substring_copy_number_vector cn;
				h->reduced_states[s].unique_substring_count_idxs_lhs(h->reduced[s].lb, cn);
				std::uint32_t i(0);
				for (auto &c : cn)
					c.string_idx = i++;
				std::size_t sum(0);
				for (auto &c : cn)
				{
					sum += c.copy_number;
					c.copy_number = sum;
				}
				fseq::create_segment_texts_task task(h->reduced_states[s], cn, h->K, seq_count, h->segment_texts[s]);
				task.execute();
			}
		}
		return 0;
	}

	double fsqref_first_pass_seconds(void *hp) { return static_cast <harness *>(hp)->first_pass_seconds; }
	std::uint32_t fsqref_K(void *hp) { return static_cast <harness *>(hp)->K; }
	std::uint64_t fsqref_dp_size(void *hp) { return static_cast <harness *>(hp)->dp.size(); }
	std::uint64_t fsqref_traceback_size(void *hp) { return static_cast <harness *>(hp)->traceback.size(); }
	std::uint64_t fsqref_reduced_size(void *hp) { return static_cast <harness *>(hp)->reduced.size(); }

	static void copy_args(fseq::segmentation_traceback_vector const &v, std::uint64_t *lb_, std::uint64_t *rb, std::uint32_t *mx, std::uint32_t *sz)
	{
		for (std::size_t i(0); i < v.size(); ++i)
		{
			lb_[i] = v[i].lb;
			rb[i] = v[i].rb;
			mx[i] = v[i].segment_max_size;
			sz[i] = v[i].segment_size;
		}
	}

	void fsqref_get_dp(void *hp, std::uint64_t *lb_, std::uint64_t *rb, std::uint32_t *mx, std::uint32_t *sz)
	{ copy_args(static_cast <harness *>(hp)->dp, lb_, rb, mx, sz); }
	void fsqref_get_traceback(void *hp, std::uint64_t *lb_, std::uint64_t *rb, std::uint32_t *mx, std::uint32_t *sz)
	{ copy_args(static_cast <harness *>(hp)->traceback, lb_, rb, mx, sz); }
	void fsqref_get_reduced(void *hp, std::uint64_t *lb_, std::uint64_t *rb, std::uint32_t *mx, std::uint32_t *sz)
	{ copy_args(static_cast <harness *>(hp)->reduced, lb_, rb, mx, sz); }

	static void copy_states(std::vector <fseq::pbwt_sample_type> const &v, std::uint32_t m, std::uint32_t *a, std::uint32_t *d)
	{
		for (std::size_t s(0); s < v.size(); ++s)
		{
			auto const &pa(v[s].input_permutation());
			auto const &pd(v[s].input_divergence());
			for (std::uint32_t i(0); i < m; ++i)
			{
				a[s * m + i] = pa[i];
				d[s * m + i] = pd[i];
			}
		}
	}

	void fsqref_get_traceback_states(void *hp, std::uint32_t *a, std::uint32_t *d)
	{ auto *h(static_cast <harness *>(hp)); copy_states(h->traceback_states, h->m, a, d); }
	void fsqref_get_reduced_states(void *hp, std::uint32_t *a, std::uint32_t *d)
	{ auto *h(static_cast <harness *>(hp)); copy_states(h->reduced_states, h->m, a, d); }

	// Segment texts of reduced segment s: K slots.  slot_off has K+1 entries, ids capacity m, copied_from K entries
	// (UINT64_MAX for an original).
	void fsqref_get_segment_texts(void *hp, std::uint64_t s, std::uint64_t *slot_off, std::uint64_t *ids, std::uint64_t *copied_from)
	{
		auto *h(static_cast <harness *>(hp));
		auto const &slots(h->segment_texts[s]);
		std::uint64_t off(0);
		for (std::size_t i(0); i < slots.size(); ++i)
		{
			slot_off[i] = off;
			for (auto const id : slots[i].sequence_indices)
				ids[off++] = id;
			copied_from[i] = slots[i].copied_from;
		}
		slot_off[slots.size()] = off;
	}

	// merge_segments_task.cc:133-176 (INTERSECTION scoring, main.cc:126): K×K weights between reduced segments s and s+1.
	void fsqref_get_pair_weights(void *hp, std::uint64_t s, std::int32_t *W)
	{
		auto *h(static_cast <harness *>(hp));
		auto const &lhs(h->segment_texts[s]);
		auto const &rhs(h->segment_texts[s + 1]);
		auto const K(lhs.size());
		for (std::size_t i(0); i < K; ++i)
			for (std::size_t j(0); j < K; ++j)
				if (! (lhs[i].is_copied() || rhs[j].is_copied()))
					W[i * K + j] = lb::set_intersection_size(
						lhs[i].sequence_indices.cbegin(), lhs[i].sequence_indices.cend(),
						rhs[j].sequence_indices.cbegin(), rhs[j].sequence_indices.cend());
		for (std::size_t i(0); i < K; ++i)
			for (std::size_t j(0); j < K; ++j)
				if (lhs[i].is_copied() || rhs[j].is_copied())
					W[i * K + j] = W[lhs[i].row_number(i) * K + rhs[j].row_number(j)];
	}

	// Short path, segmentation_sp_context.cc:21-28: distinct rows in PBWT order with run lengths.
	std::uint64_t fsqref_short_path(void *hp, std::uint32_t *row_of_run, std::uint32_t *run_len)
	{
		auto *h(static_cast <harness *>(hp));
		fseq::pbwt_context ctx(h->sequences, h->alphabet);
		ctx.prepare();
		ctx.process <>(h->n, [](){});
		fseq::substring_copy_number_vector perm;
		auto const count(ctx.unique_substring_count_idxs_lhs(0, perm));
		for (std::size_t i(0); i < perm.size(); ++i)
		{
			row_of_run[i] = perm[i].substring_idx;
			run_len[i] = perm[i].copy_number;
		}
		return count;
	}

	// rmq.hh, compiled as is: build over vals[0..count) by appending, then answer queries.
	void fsqref_rmq_queries(std::uint32_t const *vals, std::uint64_t count, std::uint64_t const *beg, std::uint64_t const *end, std::uint64_t nq, std::uint64_t *out)
	{
		fseq::segmentation_traceback_vector tb(count);
		fseq::segmentation_traceback_vector_rmq r(tb);
		for (std::uint64_t i(0); i < count; ++i)
		{
			tb[i] = fseq::segmentation_dp_arg(0, 0, vals[i], vals[i]);
			r.update(i);
		}
		for (std::uint64_t q(0); q < nq; ++q)
			out[q] = r(beg[q], end[q]);
	}
}
