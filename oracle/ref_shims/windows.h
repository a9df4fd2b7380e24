/* Minimal stand-in for <windows.h> so that the reference's scene.cpp
 * (reference scene.cpp:3, 149-156, 691-753) compiles on Linux for the oracle
 * build.  TEST INFRASTRUCTURE ONLY: no key is ever pressed. */
#pragma once
#define VK_SHIFT 0x10
#define VK_LEFT  0x25
#define VK_UP    0x26
#define VK_RIGHT 0x27
#define VK_DOWN  0x28
static inline short GetAsyncKeyState(int) { return 0; }
