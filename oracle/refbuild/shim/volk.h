/*
 * Compile-only stub, NOT a Vulkan loader: the reference's Header/Configs.h includes "volk.h"
 * solely for two constexpr enum values (Configs.h:19, :65).  Camera.cpp (pure CPU glm math)
 * includes Configs.h for CameraConfig/ClusterForwardConfig, so to compile Camera.cpp where it
 * lies under /root/reference we supply the two enum types it names.  No function is declared,
 * nothing here can create a device, and nothing in this repo links against it.
 */
#ifndef HVREF_VOLK_STUB_H
#define HVREF_VOLK_STUB_H
typedef enum VkPresentModeKHR { VK_PRESENT_MODE_FIFO_KHR = 2 } VkPresentModeKHR;
typedef enum VkFormat { VK_FORMAT_R32G32B32A32_SFLOAT = 109 } VkFormat;
#endif
