// Headless stand-in for the reference's main.cpp + openglapp.cpp (GLUT window, GL texture: unavailable on this box):
// OpenCL::init(), then the idle callback's update() (openglapp.cpp:49) until the output file is written.
#include <chrono>
#include <cstdio>
#include <iostream>

#include "OpenCLApp.h"

bool mcpt_dropin_finished();
void mcpt_dropin_shutdown();

int main() {
	try {
		const auto t0 = std::chrono::steady_clock::now();
		MCPT::OpenCL::init();
		const auto t1 = std::chrono::steady_clock::now();
		long frames = 0;
		while (!mcpt_dropin_finished()) { MCPT::OpenCL::update(); ++frames; }
		const auto t2 = std::chrono::steady_clock::now();
		std::printf("{\"init_s\": %.3f, \"update_loop_s\": %.3f, \"update_calls\": %ld}\n",
		            std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(t2 - t1).count(), frames);
		mcpt_dropin_shutdown();
	} catch (const char* msg) {
		std::cerr << "error: " << msg << std::endl;
		return 1;
	}
	return 0;
}
