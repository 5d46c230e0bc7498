// Test-infrastructure shim (oracle/_ref build only): boost::format is used by
// the reference only to build assertion messages. Not product code.
#pragma once
#include <ostream>
#include <sstream>
#include <string>
namespace boost {
	class format {
		std::string m_fmt; std::ostringstream m_args;
	public:
		explicit format(char const *f): m_fmt(f) {}
		explicit format(std::string const &f): m_fmt(f) {}
		format(format const &o): m_fmt(o.m_fmt) { m_args << o.m_args.str(); }
		template <typename T> format &operator%(T const &v) { m_args << ' ' << v; return *this; }
		std::string str() const { return m_fmt + m_args.str(); }
	};
	inline std::string str(format const &f) { return f.str(); }
	inline std::ostream &operator<<(std::ostream &os, format const &f) { return os << f.str(); }
}
