// What is left of oclbasic.cpp once the OpenCL layer is gone: the statics oclbasic.h declares (the header stays for
// its host float4 math) and entry points that the kept reference files still reference but never reach
// (thirdpartywrapper.cpp's outputPicture(cl::Buffer), treeletBVH.cpp's TreeletBVH<GPU>).
#include <string>
#include <vector>
#undef M_PI
#include "oclbasic.h"

namespace MCPT {
std::vector<cl::Platform> OpenCLBasic::platforms;
std::vector<cl::Device> OpenCLBasic::devices;
cl::Context OpenCLBasic::context;
cl::CommandQueue OpenCLBasic::queue;
bool OpenCLBasic::has_init = false;
static const char* const GONE = "the OpenCL layer was replaced by libmcpt_cuda.so";
void OpenCLBasic::init() { throw GONE; }
cl::Context& OpenCLBasic::getContext() { return context; }
cl::CommandQueue& OpenCLBasic::getQueue() { return queue; }
cl::Program OpenCLBasic::createProgramFromFileWithHeader(const std::string&, const std::string&, std::string) { throw GONE; }
cl::Kernel OpenCLBasic::createKernel(cl::Program&, const std::string&) { throw GONE; }
void OpenCLBasic::enqueueReadBuffer(const cl::Buffer&, size_t, size_t, void*) { throw GONE; }
void OpenCLBasic::enqueue1DKernelWithGroupCount(cl::Kernel&, size_t, size_t, const std::vector<cl::Event>*, cl::Event*) { throw GONE; }
void OpenCLBasic::enqueueNDRange(cl::Kernel&, cl::NDRange, cl::NDRange, const std::vector<cl::Event>*, cl::Event*) { throw GONE; }
void OpenCLBasic::readBuffer(const cl::Buffer&, void*) { throw GONE; }
float OpenCLBasic::timeCost(const cl::Event&, int) { return 0.f; }
}  // namespace MCPT
