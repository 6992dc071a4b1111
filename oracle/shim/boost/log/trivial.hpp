// TEST INFRASTRUCTURE — not product code.
// No-op stand-in for the Boost.Log headers that /root/reference/src/Logging.h:6-11 includes, so the
// reference's kernel-side sources build without linking Boost.Log. LOG(lvl) << ... compiles to a
// stream that discards everything.
#pragma once
#include <cstddef>

namespace boost { namespace log {

struct er_null_stream {
    template <typename T> er_null_stream& operator<<(const T&) { return *this; }
    er_null_stream& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};

namespace trivial {
enum severity_level { trace, debug, info, warning, error, fatal };
struct logger { static int get() { return 0; } };
}

namespace keywords {
struct er_severity_kw { int operator=(int v) const { return v; } };
static const er_severity_kw severity{};
}

template <typename T> inline int add_value(const char*, const T&) { return 0; }

namespace attributes {
struct local_clock { typedef int value_type; };
struct current_thread_id { typedef int value_type; };
}

}}  // namespace boost::log

#define BOOST_LOG_STREAM_WITH_PARAMS(logger, params) ::boost::log::er_null_stream()
#define BOOST_LOG_ATTRIBUTE_KEYWORD(name, str, type)
