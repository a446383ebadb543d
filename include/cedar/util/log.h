// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/util/log.h -- levelled logging (status, info, error, debug, memory,
// timer) selected by the "log" key of the JSON config, like the reference's
// include/cedar/util/log.h; errors are reported, never thrown.
#ifndef CEDAR_LOG_H
#define CEDAR_LOG_H

#include <iostream>
#include <sstream>
#include <string>
#include <vector>

namespace cedar { namespace log {

enum level_bits : unsigned { status_bit = 1, info_bit = 2, error_bit = 4, memory_bit = 8, debug_bit = 16, timer_bit = 32 };

unsigned &level_mask();                       // default: status | error
void set_levels(const std::vector<std::string> &names);
void set_header_msg(const std::string &msg);  // appended to the line header

class channel
{
public:
	channel(unsigned bit, const char *tag, bool to_err) : bit(bit), tag(tag), to_err(to_err) {}
	bool active() const { return (level_mask() & bit) != 0; }
	template <class T> channel &operator<<(const T &v)
	{
		if (active()) { begin(); out() << v; }
		return *this;
	}
	channel &operator<<(std::ostream &(*manip)(std::ostream &))
	{
		if (active()) {
			begin();
			out() << manip;
			// std::endl / std::flush end the current line: the next insertion prints the header again
			fresh = true;
		}
		return *this;
	}
private:
	std::ostream &out() { return to_err ? std::cerr : std::cout; }
	void begin();
	unsigned bit;
	const char *tag;
	bool to_err;
	bool fresh = true;
};

extern channel status, info, error, memory, debug, timer;

}}

#endif
