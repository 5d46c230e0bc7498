// The seam between the host driver and whatever computes the segmentation (SURVEY.md 8b, S1 / S1'): what
// founder_sequences::segmentation_container holds, as plain arrays.  founder_sequences_b200 links engine_fsq.cc (the C
// ABI of libfsq_b200.so: include/fsq.h).  The test harness links oracle/engine_ref.cc instead (the reference's own
// arithmetic compiled in place) and produces the files the product's are compared with byte for byte.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace fsq_host {

struct segmentation
{
	bool short_path{false};                 // n < 2L: row_of_run / run_len are filled, the rest is empty
	std::vector <std::uint32_t> row_of_run, run_len;
	std::uint32_t K{};                      // container.max_segment_size
	std::vector <std::uint64_t> lb, rb;     // container.reduced_traceback[i].{lb, rb}
	std::vector <std::uint32_t> size;       //                                .segment_size
	std::vector <std::uint32_t> a, d;       // container.reduced_pbwt_samples[i].input_permutation() / input_divergence(), count x m
};

enum engine_status { ENGINE_OK = 0, ENGINE_NOT_REDUCIBLE = 1, ENGINE_ERROR = 2 };

// rows: m pointers to n bytes.  comp: the alphabet's char -> comp table (consecutive_alphabet.hh:259-264).
engine_status run_engine(std::uint8_t const *const *rows, std::uint32_t m, std::uint64_t n, std::uint8_t const comp[256], std::uint64_t segment_length_bound,
	segmentation &out, std::string &message);

char const *engine_name();

} // namespace fsq_host
