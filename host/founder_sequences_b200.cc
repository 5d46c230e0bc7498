// founder_sequences_b200 — a command-line compatible host driver around the C ABI of libfsq_b200.so.
//
// The reference's own host (founder-sequences/main.cc, generate_context.cc, join_context.cc) needs clang + Blocks,
// libdispatch, Boost, gengetopt and LEMON, none of which this image has.  This driver takes the same
// options (founder-sequences/cmdline.ggo:12-29), reads the same inputs, writes the same two files, and keeps the
// division of labour north_star asks for: segmentation on the GPU through include/fsq.h, joining and writing on the
// host with the reference's OWN joiner and writer code compiled in place from the reference tree, never copied:
//   founder-sequences/greedy_matcher.cc      greedy_matcher::match / output_segments
//   founder-sequences/segmentation_dp_arg.cc the segments.txt writers
//   founder-sequences/segment_text.cc
//   lib/libbio pbwt_context::unique_substring_count_idxs_lhs (header)   runs and copy numbers
// What is restated here, because the originals sit in files that include LEMON or use Blocks (each cites its lines):
//   the controller's load / check / alphabet steps          generate_context.cc:64-158
//   copy numbers per segment, random and PBWT-order joining  join_context.cc:51-160, 270-330
//   the row-wise founder writer                               join_context.cc:333-357
//   the short path's two writers                              segmentation_sp_context.cc:31-47
// Bipartite joining ends in LEMON 1.3.1's MaxWeightedPerfectMatching, which is not in the reference tree (SURVEY.md
// F8): asked for, it is refused with that explanation rather than answered with some other matching.
#include "engine.hh"

#include <founder_sequences/founder_sequences.hh>
#include <founder_sequences/greedy_matcher.hh>
#include <founder_sequences/segmentation_dp_arg.hh>
#include <founder_sequences/substring_copy_number.hh>
#include <libbio/bits.hh>

#include <sys/mman.h>
#include <sys/stat.h>
#include <fcntl.h>
#include <getopt.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <random>

namespace fseq = founder_sequences;
namespace lb = libbio;

// libbio::dispatch_ptr pulls these two symbols in; nothing from libdispatch is used (the queues stay null).
extern "C" void dispatch_retain(dispatch_object_t) {}
extern "C" void dispatch_release(dispatch_object_t) {}

namespace {

// ---- input (lib/libbio/src/sequence_reader.cc:44-53: one mmap'd file per row; FASTA: one record per row) -------------
struct input_rows
{
	std::vector <std::pair <void *, std::size_t>> maps;
	std::vector <std::vector <std::uint8_t>> owned;
	fseq::sequence_vector sequences;

	~input_rows() { for (auto &mp : maps) if (mp.first && mp.second) munmap(mp.first, mp.second); }

	bool add_file(std::string const &path)
	{
		int const fd(open(path.c_str(), O_RDONLY));
		if (fd < 0) { std::cerr << "Unable to open " << path << ": " << strerror(errno) << std::endl; return false; }
		struct stat sb;
		if (fstat(fd, &sb)) { close(fd); return false; }
		void *p(nullptr);
		if (sb.st_size)
		{
			p = mmap(nullptr, sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
			if (MAP_FAILED == p) { std::cerr << "Unable to map " << path << std::endl; close(fd); return false; }
		}
		close(fd);
		maps.emplace_back(p, sb.st_size);
		sequences.emplace_back(static_cast <std::uint8_t const *>(p), sb.st_size);
		return true;
	}

	bool read_list(char const *path)
	{
		std::ifstream file;
		std::istream *in(&std::cin);
		if (path && !('-' == path[0] && '\0' == path[1]))
		{
			file.open(path);
			if (!file) { std::cerr << "Unable to open " << path << std::endl; return false; }
			in = &file;
		}
		std::string line;
		while (std::getline(*in, line))
			if (!add_file(line)) return false;
		return true;
	}

	// the subset of FASTA the reference's Ragel reader accepts for this tool: '>' header lines, sequence lines joined
	bool read_fasta(char const *path)
	{
		std::ifstream file(path);
		if (!file) { std::cerr << "Unable to open " << path << std::endl; return false; }
		std::string line;
		while (std::getline(file, line))
		{
			if (!line.empty() && '\r' == line.back()) line.pop_back();
			if (!line.empty() && '>' == line[0]) { owned.emplace_back(); continue; }
			if (owned.empty()) continue;
			owned.back().insert(owned.back().end(), line.begin(), line.end());
		}
		for (auto const &v : owned) sequences.emplace_back(v.data(), v.size());
		return true;
	}
};

// pbwt_context's members are protected with no setters (lib/libbio/include/libbio/pbwt_context.hh:135-160): a derived
// class fills what the joiners read -- input_permutation(), input_divergence(), sequence_idx().
struct filled_sample final : public fseq::pbwt_sample_type
{
	filled_sample(std::uint32_t const *a, std::uint32_t const *d, std::uint32_t m, std::size_t rb)
	{
		m_input_permutation = string_index_vector(m, 0);
		m_input_divergence = character_index_vector(m, 0);
		for (std::uint32_t i(0); i < m; ++i) { m_input_permutation[i] = a[i]; m_input_divergence[i] = d[i]; }
		m_sequence_idx = rb;
	}
};

struct options
{
	char const *input{};
	bool fasta{false};
	char const *out_founders{};
	char const *out_segments{};
	long segment_length{-1};
	fseq::segment_joining joining{fseq::segment_joining::BIPARTITE_MATCHING};
	long sample_rate{4};
	long seed{0};
	bool print_invocation{false};
};

void usage()
{
	std::cerr <<
		"founder_sequences_b200 --input=input-list.txt --segment-length-bound=SIZE [--input-format=FASTA|list-file]\n"
		"    [--output-founders=PATH] [--output-segments=PATH] [--segment-joining=bipartite-matching|greedy|random|pbwt-order]\n"
		"    [--pbwt-sample-rate=q] [--random-seed=N] [--single-threaded] [--print-invocation]\n"
		"The options of founder_sequences (tsnorri/founder-sequences); segmentation on the GPU through libfsq_b200.so.\n";
}

// The joiners' view of the run: what join_context is to the reference's matchers (join_context.hh:26-107).
class joiner final : public fseq::greedy_matcher_delegate
{
public:
	fseq::sequence_vector const				*m_sequences{};
	std::vector <fseq::pbwt_sample_type>	m_samples;
	fseq::segmentation_traceback_vector		m_traceback;
	std::uint32_t							m_max_segment_size{};
	fseq::substring_copy_number_matrix		m_substring_copy_numbers;
	fseq::permutation_matrix				m_permutations;
	std::unique_ptr <fseq::matcher>			m_matcher;
	std::uint32_t							m_permutation_max{};
	std::uint8_t							m_permutation_bits_needed{};
	std::ostream							*m_founders{};
	bool									m_finished{false};

	lb::dispatch_ptr <dispatch_queue_t> producer_queue() const override { return {}; }
	lb::dispatch_ptr <dispatch_queue_t> consumer_queue() const override { return {}; }
	std::vector <fseq::pbwt_sample_type> const &pbwt_samples() const override { return m_samples; }
	fseq::segmentation_traceback_vector const &reduced_traceback() const override { return m_traceback; }
	fseq::permutation_matrix &permutations() override { return m_permutations; }
	fseq::bipartite_set_scoring bipartite_set_scoring_method() const override { return fseq::bipartite_set_scoring::INTERSECTION; }
	std::uint32_t sequence_count() const override { return m_sequences->size(); }
	std::uint32_t max_segment_size() const override { return m_max_segment_size; }

	void matcher_did_finish(fseq::greedy_matcher &) override
	{
		output_in_permutation_order();
		m_finished = true;
	}

	// join_context.cc:51-160 (everything but GREEDY fills m_substring_copy_numbers)
	void count_substrings(fseq::segment_joining const seg_joining)
	{
		m_substring_copy_numbers.clear();
		if (fseq::segment_joining::GREEDY == seg_joining) return;
		m_substring_copy_numbers.resize(m_traceback.size());
		for (std::size_t s(0); s < m_traceback.size(); ++s)
		{
			auto const &dp_arg(m_traceback[s]);
			auto const &sample(m_samples[s]);
			auto &substring_cn(m_substring_copy_numbers[s]);
			auto const substring_count(sample.unique_substring_count_idxs_lhs(dp_arg.lb, substring_cn));
			{
				std::uint32_t i(0);
				for (auto &cn : substring_cn) cn.string_idx = i++;
			}
			if (fseq::segment_joining::BIPARTITE_MATCHING != seg_joining)
			{
				std::sort(substring_cn.begin(), substring_cn.end());                     // :87, by count
				auto const empty_slots(m_max_segment_size - substring_count);
				std::size_t remaining_slots(empty_slots);
				for (auto it(substring_cn.rbegin()); it != substring_cn.rend(); ++it)   // :92-97
				{
					auto const addition(std::min <std::size_t>(remaining_slots, std::ceil(1.0 * it->copy_number / substring_count * empty_slots)));
					it->copy_number = 1 + addition;
					remaining_slots -= addition;
				}
				while (remaining_slots)                                                  // :100-110
					for (auto it(substring_cn.rbegin()); it != substring_cn.rend() && remaining_slots; ++it)
					{
						++it->copy_number;
						--remaining_slots;
					}
				if (fseq::segment_joining::PBWT_ORDER == seg_joining)                    // :116-121
					std::sort(substring_cn.begin(), substring_cn.end(), [](auto const &l, auto const &r){ return l.string_idx < r.string_idx; });
			}
			std::size_t sum(0);                                                          // make_cumulative_sum
			for (auto &cn : substring_cn) { sum += cn.copy_number; cn.copy_number = sum; }
		}
	}

	// join_context.cc:191-208
	void init_permutations()
	{
		auto const seq_count(sequence_count());
		m_permutation_bits_needed = lb::bits::highest_bit_set(seq_count);
		m_permutation_max = (1 << m_permutation_bits_needed) - 1;
		m_permutations.clear();
		m_permutations.resize(m_traceback.size());
		for (auto &permutation : m_permutations)
		{
			sdsl::int_vector <0> temp(m_max_segment_size, 0, m_permutation_bits_needed);
			permutation = std::move(temp);
		}
	}

	// join_context.cc:333-357; gaps for the "no string" marker as in output_gaps
	void output_in_permutation_order() const
	{
		auto &stream(*m_founders);
		for (std::size_t row(0); row < m_max_segment_size; ++row)
		{
			for (std::size_t s(0); s < m_traceback.size(); ++s)
			{
				auto const &dp_arg(m_traceback[s]);
				auto const substring_idx(m_permutations[s][row]);
				auto const len(dp_arg.rb - dp_arg.lb);
				if (m_permutation_max == substring_idx)
					for (std::size_t i(0); i < len; ++i) stream << '-';
				else
					stream.write(reinterpret_cast <char const *>((*m_sequences)[substring_idx].data() + dp_arg.lb), len);
			}
			stream << '\n';
		}
		stream << std::flush;
	}

	// join_context.cc:270-296
	void join_random_order_and_output(std::uint_fast32_t const seed)
	{
		init_permutations();
		std::mt19937 urbg(seed);
		for (std::size_t s(0); s < m_traceback.size(); ++s)
		{
			auto const &cn_vector(m_substring_copy_numbers[s]);
			auto &permutation(m_permutations[s]);
			auto const begin(permutation.begin());
			auto it(permutation.begin());
			for (auto const &cn : cn_vector)
			{
				auto end(begin + cn.copy_number);
				std::fill(it, end, cn.substring_idx);
				it = end;
			}
			std::shuffle(permutation.begin(), permutation.end(), urbg);
		}
		output_in_permutation_order();
	}

	// join_context.cc:299-330
	void join_pbwt_order_and_output() const
	{
		auto &stream(*m_founders);
		std::vector <fseq::substring_copy_number_vector::const_iterator> its(m_substring_copy_numbers.size());
		for (std::size_t s(0); s < its.size(); ++s) its[s] = m_substring_copy_numbers[s].cbegin();
		for (std::size_t row(0); row < m_max_segment_size; ++row)
		{
			for (std::size_t s(0); s < m_traceback.size(); ++s)
			{
				auto &it(its[s]);
				auto const &dp_arg(m_traceback[s]);
				if (row == it->copy_number) ++it;
				stream.write(reinterpret_cast <char const *>((*m_sequences)[it->substring_idx].data() + dp_arg.lb), dp_arg.rb - dp_arg.lb);
			}
			stream << '\n';
		}
		stream << std::flush;
	}
};

} // namespace


int main(int argc, char **argv)
{
	options opt;
	static option const long_options[] = {
		{"input", required_argument, nullptr, 'i'}, {"input-format", required_argument, nullptr, 'f'},
		{"output-segments", required_argument, nullptr, 'e'}, {"output-founders", required_argument, nullptr, 'o'},
		{"segment-length-bound", required_argument, nullptr, 's'}, {"segment-joining", required_argument, nullptr, 'j'},
		{"pbwt-sample-rate", required_argument, nullptr, 'm'}, {"random-seed", required_argument, nullptr, 1},
		{"single-threaded", no_argument, nullptr, 2}, {"print-invocation", no_argument, nullptr, 3},
		{"help", no_argument, nullptr, 'h'}, {nullptr, 0, nullptr, 0}};
	int c;
	while (-1 != (c = getopt_long(argc, argv, "i:f:e:o:s:j:m:h", long_options, nullptr)))
	{
		switch (c)
		{
			case 'i': opt.input = optarg; break;
			case 'f':
				if (0 == strcmp(optarg, "FASTA")) opt.fasta = true;
				else if (0 == strcmp(optarg, "list-file")) opt.fasta = false;
				else { std::cerr << "Unexpected value for input_format" << std::endl; return EXIT_FAILURE; }
				break;
			case 'e': opt.out_segments = optarg; break;
			case 'o': opt.out_founders = optarg; break;
			case 's': opt.segment_length = atol(optarg); break;
			case 'j':
				if (0 == strcmp(optarg, "bipartite-matching")) opt.joining = fseq::segment_joining::BIPARTITE_MATCHING;
				else if (0 == strcmp(optarg, "greedy")) opt.joining = fseq::segment_joining::GREEDY;
				else if (0 == strcmp(optarg, "random")) opt.joining = fseq::segment_joining::RANDOM;
				else if (0 == strcmp(optarg, "pbwt-order")) opt.joining = fseq::segment_joining::PBWT_ORDER;
				else { std::cerr << "Unexpected value for segment_joining." << std::endl; return EXIT_FAILURE; }
				break;
			case 'm': opt.sample_rate = atol(optarg); break;
			case 1: opt.seed = atol(optarg); break;
			case 2: break;                       // one host thread is all this driver uses
			case 3: opt.print_invocation = true; break;
			case 'h': usage(); return EXIT_SUCCESS;
			default: usage(); return EXIT_FAILURE;
		}
	}
	std::ios_base::sync_with_stdio(false);
	if (opt.print_invocation)
	{
		std::cerr << "Invocation:";
		for (int i(0); i < argc; ++i) std::cerr << ' ' << argv[i];
		std::cerr << std::endl;
	}
	// main.cc:91-115
	if (!opt.input) { usage(); return EXIT_FAILURE; }
	if (opt.segment_length < 0 && -1 == opt.segment_length) { std::cerr << "Segment length bound needs to be specified when generating a segmentation." << std::endl; return EXIT_FAILURE; }
	if (opt.segment_length <= 0) { std::cerr << "Segment length bound must be positive." << std::endl; return EXIT_FAILURE; }
	if (!(0 <= opt.seed && (unsigned long) opt.seed <= std::numeric_limits <std::uint_fast32_t>::max())) { std::cerr << "Random seed out of bounds." << std::endl; return EXIT_FAILURE; }
	if (opt.sample_rate <= 0) { std::cerr << "PBWT sample rate multiplier must be non-negative." << std::endl; return EXIT_FAILURE; }
	if (fseq::segment_joining::BIPARTITE_MATCHING == opt.joining)
	{
		std::cerr << "Bipartite joining ends in LEMON 1.3.1's MaxWeightedPerfectMatching, which the reference downloads at build time and which is not "
			"available here; its choice among equal-weight matchings decides the founders, so no stand-in is offered. Use --segment-joining=greedy, random or pbwt-order."
			<< std::endl;
		return EXIT_FAILURE;
	}

	// generate_context.cc:64-106
	input_rows in;
	std::cerr << "Loading the input…" << std::flush;
	if (!(opt.fasta ? in.read_fasta(opt.input) : in.read_list(opt.input))) return EXIT_FAILURE;
	auto const &sequences(in.sequences);
	if (0 == sequences.size()) { std::cerr << "\nThe input file contained no sequences." << std::endl; return EXIT_SUCCESS; }
	auto const seq_length(sequences.front().size());
	std::cerr << " length: " << seq_length << std::endl;
	{
		bool stop(false);
		for (std::size_t i(1); i < sequences.size(); ++i)
			if (sequences[i].size() != seq_length)
			{
				stop = true;
				std::cerr << "The length of the sequence at index " << i << " was " << sequences[i].size() << " while that of the first one was " << seq_length << '.' << std::endl;
			}
		if (stop) return EXIT_FAILURE;
	}

	// generate_context.cc:134-147: the reference's own alphabet builder (comps in byte order, NUL = 0)
	std::cerr << "Generating a compressed alphabet…" << std::endl;
	fseq::alphabet_type alphabet;
	{
		lb::consecutive_alphabet_as_builder <std::uint8_t> builder;
		builder.init();
		for (auto const &vec : sequences) builder.prepare(vec);
		builder.compress();
		using std::swap;
		swap(alphabet, builder.alphabet());
	}
	std::uint8_t comp[256];
	for (unsigned ch(0); ch < 256; ++ch) comp[ch] = 0;
	for (auto const &vec : sequences) for (auto const ch : vec) comp[ch] = alphabet.char_to_comp(ch);

	std::ofstream founders_file, segments_file;
	std::ostream *founders(&std::cout), *segments(nullptr);
	if (opt.out_founders) { founders_file.open(opt.out_founders, std::ios::binary); if (!founders_file) { std::cerr << "Unable to open " << opt.out_founders << std::endl; return EXIT_FAILURE; } founders = &founders_file; }
	if (opt.out_segments) { segments_file.open(opt.out_segments, std::ios::binary); if (!segments_file) { std::cerr << "Unable to open " << opt.out_segments << std::endl; return EXIT_FAILURE; } segments = &segments_file; }

	// generate_context.cc:379-390: the segmentation (engine: host/engine.hh)
	std::cerr << "Calculating the segmentation (" << fsq_host::engine_name() << ")…" << std::endl;
	std::vector <std::uint8_t const *> row_ptrs(sequences.size());
	for (std::size_t i(0); i < sequences.size(); ++i) row_ptrs[i] = sequences[i].data();
	fsq_host::segmentation seg;
	std::string message;
	auto const status(fsq_host::run_engine(row_ptrs.data(), sequences.size(), seq_length, comp, opt.segment_length, seg, message));
	if (fsq_host::ENGINE_NOT_REDUCIBLE == status)
	{
		// generate_context.cc:192-200
		std::cerr << "Unable to reduce the number of sequences; the maximum segment size is equal to the number of input sequences." << std::endl;
		return EXIT_FAILURE;
	}
	if (fsq_host::ENGINE_OK != status) { std::cerr << "Segmentation failed: " << message << std::endl; return EXIT_FAILURE; }

	if (seg.short_path)
	{
		// segmentation_sp_context.cc:31-47
		std::cerr << "Outputting…" << std::endl;
		for (auto const row : seg.row_of_run)
		{
			founders->write(reinterpret_cast <char const *>(sequences[row].data()), sequences[row].size());
			*founders << '\n';
		}
		*founders << std::flush;
		if (segments)
		{
			*segments << "SEQUENCE" "\n";
			for (auto const len : seg.run_len) *segments << len << '\n';
			*segments << std::flush;
		}
		std::cerr << "Done." << std::endl;
		return EXIT_SUCCESS;
	}

	std::cerr << "Maximum segment size: " << seg.K << "; reduced segments: " << seg.lb.size() << std::endl;
	joiner J;
	J.m_sequences = &sequences;
	J.m_max_segment_size = seg.K;
	J.m_founders = founders;
	std::uint32_t const m(sequences.size());
	for (std::size_t s(0); s < seg.lb.size(); ++s)
	{
		J.m_traceback.emplace_back(seg.lb[s], seg.rb[s], seg.size[s]);       // segmentation_dp_arg.hh:27-36 (three arguments, as find_segments_greedy emits)
		J.m_samples.emplace_back(filled_sample(seg.a.data() + s * m, seg.d.data() + s * m, m, seg.rb[s]));
	}

	std::cerr << "Joining the segments…" << std::endl;
	J.count_substrings(opt.joining);
	switch (opt.joining)
	{
		case fseq::segment_joining::GREEDY:
			J.init_permutations();
			J.m_matcher.reset(new fseq::greedy_matcher(J, J.m_substring_copy_numbers, J.m_permutation_max, J.m_permutation_bits_needed));
			J.m_matcher->match();
			if (!J.m_finished) { std::cerr << "The greedy matcher did not finish." << std::endl; return EXIT_FAILURE; }
			if (segments) J.m_matcher->output_segments(*segments, sequences);
			break;
		case fseq::segment_joining::RANDOM:
			J.join_random_order_and_output(opt.seed);
			if (segments) fseq::output_segments(*segments, J.m_traceback, J.m_substring_copy_numbers, sequences);
			break;
		case fseq::segment_joining::PBWT_ORDER:
			J.join_pbwt_order_and_output();
			if (segments) fseq::output_segments(*segments, J.m_traceback, J.m_substring_copy_numbers, sequences);
			break;
		default:
			return EXIT_FAILURE;
	}
	std::cerr << "Done." << std::endl;
	return EXIT_SUCCESS;
}
