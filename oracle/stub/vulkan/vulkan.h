/* Test-infrastructure stub: lets the reference's headers see Vulkan enum/handle
 * names without a Vulkan SDK. Wraps the glad header vendored by the reference. */
#pragma once
#include "glad/vulkan.h"
typedef struct VkDebugUtilsMessengerEXT_T* VkDebugUtilsMessengerEXT;
