// TEST INFRASTRUCTURE ONLY -- boost::timer::cpu_timer (wall / user / system nanoseconds).
#pragma once
#include <chrono>
#include <cstdint>
#include <ctime>
#include <string>
namespace boost { namespace timer {
using nanosecond_type = std::int_least64_t;
struct cpu_times { nanosecond_type wall, user, system; void clear() { wall = user = system = 0; } };
class cpu_timer {
public:
    cpu_timer() { start(); }
    void start() { m_stopped = false; m_acc.clear(); m_t0 = now(); }
    void stop() { if (!m_stopped) { add(); m_stopped = true; } }
    void resume() { if (m_stopped) { m_stopped = false; m_t0 = now(); } }
    bool is_stopped() const { return m_stopped; }
    cpu_times elapsed() const { cpu_times t = m_acc; if (!m_stopped) { cpu_times n = now(); t.wall += n.wall - m_t0.wall; t.user += n.user - m_t0.user; t.system += n.system - m_t0.system; } return t; }
    std::string format(short = 6, const std::string& = std::string()) const { cpu_times t = elapsed(); return std::to_string(double(t.wall) * 1e-9) + "s wall, " + std::to_string(double(t.user) * 1e-9) + "s user\n"; }
private:
    static cpu_times now() {
        cpu_times t; t.wall = std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
        t.user = nanosecond_type(double(std::clock()) / CLOCKS_PER_SEC * 1e9); t.system = 0; return t;
    }
    void add() { cpu_times n = now(); m_acc.wall += n.wall - m_t0.wall; m_acc.user += n.user - m_t0.user; m_acc.system += n.system - m_t0.system; }
    cpu_times m_acc{0, 0, 0}, m_t0{0, 0, 0}; bool m_stopped = true;
};
}}
