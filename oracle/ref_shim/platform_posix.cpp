// TEST INFRASTRUCTURE ONLY (oracle). Not part of the shipped product path.
//
// POSIX stand-in for the seven Engine::Platform members that the reference's
// mesh-import translation units link against (declared in the reference at
// engine/src/platform/platform.hpp:11-12,26-30; the only implementation the
// reference ships is Win32, engine/src/platform/platform_Win32.cpp:148-171,355-373).
// It lets oracle/Makefile compile the reference's own sources, where they lie,
// into oracle/_ref/libvkref.so.
#include "platform/platform.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace Engine {

void Platform::ConsoleWriteError(std::string& text, u8) { std::fputs(text.c_str(), stderr); }
void Platform::ConsoleWrite(std::string& text, u8) { std::fputs(text.c_str(), stdout); }
void Platform::ZrMemory(void* block, u64 size) { std::memset(block, 0, size); }
void Platform::CpMemory(void* dest, const void* source, u64 size) { std::memcpy(dest, source, size); }
void Platform::FrMemory(void* block) { std::free(block); }
void* Platform::AllocMemory(u64 size) { return std::malloc(size); }
void Platform::SetMemory(void* block, i32 data, u64 size) { std::memset(block, data, size); }

}  // namespace Engine
