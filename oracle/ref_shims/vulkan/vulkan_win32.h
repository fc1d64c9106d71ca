/* Type-only stand-in for <vulkan/vulkan_win32.h> (included by the reference's include/vkEngine.h:12). */
#ifndef TPT_REF_SHIM_VULKAN_WIN32_H
#define TPT_REF_SHIM_VULKAN_WIN32_H
#include "vulkan.h"
#endif
