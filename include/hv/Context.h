// hv/Context.h -- what replaces VulkanContext / VkCommandBuffer / VulkanBuffer on this path.
//   hv::Context       device ordinal + stream + the 2-deep frame ring
//                     (VulkanContext::GetFrameIndex / IncrementFrameIndex, Source/Vulkan/VulkanContext.cpp:704-717)
//   hv::DeviceBuffer  VulkanBuffer's CreateBuffer / UploadBufferData / UploadOffsetBufferData /
//                     DownloadBufferData / Destroy (Header/Vulkan/VulkanBuffer.h:31-91) over plain device memory
// The aliases at the bottom let the reference's signatures compile unchanged:
//   FillCommandBuffer(VulkanContext& ctx, VkCommandBuffer commandBuffer)  ->  (hv::Context&, cudaStream_t as void*)
// "Barrier = stream order": everything recorded on one stream executes in order, which is what the
// reference's vkCmdPipelineBarrier2 buffer barriers enforce.
#ifndef HV_CONTEXT_H
#define HV_CONTEXT_H

#include <cstdio>
#include <stdexcept>
#include <string>

#include "../hvcf.h"
#include "Types.h"

namespace hv
{
// VK_CHECK prints and continues (Source/Vulkan/VulkanCheck.cpp:3-9); this one prints and THROWS:
// the C++ layer never continues silently after a failed call.
inline void Check(int status, const char* what)
{
	if (status != HVCF_OK)
	{
		std::string msg = std::string(what) + ": " + hvcf_last_error_string();
		std::fprintf(stderr, "hvcf error %d in %s\n", status, msg.c_str());
		throw std::runtime_error(msg);
	}
}

class Context
{
public:
	explicit Context(int device = -1, void* stream = nullptr) : device_(device), stream_(stream) {}
	uint32_t GetFrameIndex() const { return frameIndex_; }
	void IncrementFrameIndex() { frameIndex_ = (frameIndex_ + 1u) % AppConfig::FrameCount; }
	int GetDeviceOrdinal() const { return device_; }
	void* GetStream() const { return stream_; } // cudaStream_t
	void WaitIdle() const { Check(hvcf_stream_synchronize(stream_), "hvcf_stream_synchronize"); }

private:
	int device_;
	void* stream_;
	uint32_t frameIndex_ = 0;
};

class DeviceBuffer
{
public:
	void* buffer_ = nullptr; // device pointer (VulkanBuffer::buffer_)
	size_t size_ = 0;

	void CreateBuffer(Context& ctx, size_t size)
	{
		Destroy();
		Check(hvcf_device_alloc(ctx.GetDeviceOrdinal(), size, &buffer_), "hvcf_device_alloc");
		size_ = size;
		owned_ = true;
	}
	// non-owning view of a buffer that lives inside an hvcf_ctx
	void Wrap(void* devicePtr, size_t size)
	{
		Destroy();
		buffer_ = devicePtr;
		size_ = size;
		owned_ = false;
	}
	void UploadBufferData(Context& ctx, const void* data, size_t dataSize) { UploadOffsetBufferData(ctx, data, 0, dataSize); }
	void UploadOffsetBufferData(Context& ctx, const void* data, size_t offset, size_t dataSize)
	{
		if (offset + dataSize > size_)
			throw std::out_of_range("DeviceBuffer::UploadOffsetBufferData past the end");
		Check(hvcf_device_upload(buffer_, offset, data, dataSize, ctx.GetStream()), "hvcf_device_upload");
	}
	void DownloadBufferData(Context& ctx, void* outData, size_t dataSize)
	{
		if (dataSize > size_)
			throw std::out_of_range("DeviceBuffer::DownloadBufferData past the end");
		Check(hvcf_device_download(outData, buffer_, 0, dataSize, ctx.GetStream()), "hvcf_device_download");
	}
	void Destroy()
	{
		if (buffer_ && owned_)
			hvcf_device_free(buffer_);
		buffer_ = nullptr;
		size_ = 0;
		owned_ = false;
	}

private:
	bool owned_ = false;
};
} // namespace hv

using VulkanContext = hv::Context;
using VulkanBuffer = hv::DeviceBuffer;
using VkCommandBuffer = void*; // a cudaStream_t

#endif
