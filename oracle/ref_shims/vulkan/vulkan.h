/* Type-only stand-in for <vulkan/vulkan.h>: just enough declarations for the reference's
 * include/vkEngine.h and include/GLFW/glfw3.h to parse.  No Vulkan function exists or is called: there is no
 * graphics stack on this box; presentation is replaced by a headless framebuffer (oracle/ref_headless.cu). */
#ifndef TPT_REF_SHIM_VULKAN_H
#define TPT_REF_SHIM_VULKAN_H
#include <stdint.h>
#define VK_VERSION_1_0 1
#define VKAPI_PTR
#define VK_DEFINE_SHIM_HANDLE(name) typedef struct name##_T* name;
VK_DEFINE_SHIM_HANDLE(VkInstance)
VK_DEFINE_SHIM_HANDLE(VkPhysicalDevice)
VK_DEFINE_SHIM_HANDLE(VkDevice)
VK_DEFINE_SHIM_HANDLE(VkQueue)
VK_DEFINE_SHIM_HANDLE(VkCommandBuffer)
VK_DEFINE_SHIM_HANDLE(VkDebugUtilsMessengerEXT)
VK_DEFINE_SHIM_HANDLE(VkSurfaceKHR)
VK_DEFINE_SHIM_HANDLE(VkSwapchainKHR)
VK_DEFINE_SHIM_HANDLE(VkImage)
VK_DEFINE_SHIM_HANDLE(VkImageView)
VK_DEFINE_SHIM_HANDLE(VkDeviceMemory)
VK_DEFINE_SHIM_HANDLE(VkCommandPool)
VK_DEFINE_SHIM_HANDLE(VkRenderPass)
VK_DEFINE_SHIM_HANDLE(VkFramebuffer)
VK_DEFINE_SHIM_HANDLE(VkSemaphore)
VK_DEFINE_SHIM_HANDLE(VkFence)
typedef uint64_t VkDeviceSize;
typedef uint32_t VkFlags;
typedef uint32_t VkBool32;
typedef VkFlags VkCommandPoolCreateFlags;
typedef int VkResult;
typedef int VkFormat;
typedef int VkImageLayout;
typedef int VkCommandBufferLevel;
typedef int VkExternalMemoryHandleTypeFlagBits;
/* The reference hard-codes `VkExtent2D m_windowExtent{ 1920, 1080 }` (include/vkEngine.h:24) and takes the render
 * resolution from it (src/path_tracer.cu:479-480).  The stand-in's constructor lets the headless driver override the
 * extent at run time without touching the reference's header. */
extern uint32_t g_tpt_ref_width, g_tpt_ref_height;
#ifdef __cplusplus
struct VkExtent2D {
    uint32_t width, height;
    VkExtent2D(uint32_t w, uint32_t h) : width(g_tpt_ref_width ? g_tpt_ref_width : w), height(g_tpt_ref_height ? g_tpt_ref_height : h) {}
};
#else
typedef struct VkExtent2D { uint32_t width; uint32_t height; } VkExtent2D;
#endif
typedef struct VkAllocationCallbacks VkAllocationCallbacks;
typedef struct VkCommandPoolCreateInfo { int unused; } VkCommandPoolCreateInfo;
typedef struct VkCommandBufferAllocateInfo { int unused; } VkCommandBufferAllocateInfo;
typedef void (VKAPI_PTR *PFN_vkVoidFunction)(void);
#define VK_COMMAND_BUFFER_LEVEL_PRIMARY 0
#endif
