/* MSVC <intrin.h> stand-in so the reference headers (vectorMath.hpp:20, voxCommon.hpp:113-118,
 * hipUtil.hpp:6) compile with g++.  Test infrastructure only (oracle/_ref build). */
#pragma once
#include <x86intrin.h>
static inline unsigned char _BitScanForward( unsigned long* index, unsigned long mask )
{
	if( mask == 0 ) return 0;
	*index = (unsigned long)__builtin_ctzl( mask );
	return 1;
}
#ifndef __popcnt
#define __popcnt( x ) __builtin_popcount( ( x ) )
#endif
#ifndef __popcnt64
#define __popcnt64( x ) __builtin_popcountll( ( x ) )
#endif
#ifndef __debugbreak
#define __debugbreak() __builtin_trap()
#endif
