/* Type-only stand-in for <Windows.h> (included by the reference's include/vkEngine.h:11). */
#ifndef TPT_REF_SHIM_WINDOWS_H
#define TPT_REF_SHIM_WINDOWS_H
typedef void* HANDLE;
#endif
