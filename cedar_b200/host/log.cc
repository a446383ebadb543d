// log.cc -- levelled logging and timing regions (cedar/util/log.h, timer.h)
#include <chrono>
#include <ctime>
#include <map>

#include <cedar/util/log.h>
#include <cedar/util/timer.h>

namespace cedar { namespace log {

static std::string g_header_msg;

unsigned &level_mask()
{
	static unsigned mask = status_bit | error_bit;
	return mask;
}

void set_levels(const std::vector<std::string> &names)
{
	unsigned m = 0;
	for (auto &n : names) {
		if (n == "status") m |= status_bit;
		else if (n == "info") m |= info_bit;
		else if (n == "error") m |= error_bit;
		else if (n == "memory") m |= memory_bit;
		else if (n == "debug") m |= debug_bit;
		else if (n == "timer") m |= timer_bit;
	}
	level_mask() = m;
}

void set_header_msg(const std::string &msg) { g_header_msg = msg; }

void channel::begin()
{
	if (!fresh) return;
	fresh = false;
	out() << "[Cedar <" << tag << ">" << g_header_msg << "] ";
}

channel status(status_bit, "status", false), info(info_bit, "info", false), error(error_bit, "error", true),
    memory(memory_bit, "memory", false), debug(debug_bit, "debug", false), timer(timer_bit, "timer", false);

}

static std::map<std::string, std::chrono::steady_clock::time_point> g_open;

void timer_begin(const std::string &label)
{
	if (log::timer.active()) g_open[label] = std::chrono::steady_clock::now();
}
void timer_end(const std::string &label)
{
	if (!log::timer.active()) return;
	auto it = g_open.find(label);
	if (it == g_open.end()) return;
	double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - it->second).count();
	log::timer << label << " (host enqueue): " << s << " s" << std::endl;
	g_open.erase(it);
}
void timer_up() {}
void timer_down() {}

}
