// TEST INFRASTRUCTURE — headless replacement for the reference's Window.h (the original
// opens an SDL2 + OpenGL window; no graphics stack exists here). It keeps the part the
// hot path uses: the CPU frame buffer, the 32x32 tile grid (Window.h:32-33,
// Window.cpp:11-12) and the float -> 0x00RRGGBB quantisation of Window::plot
// (Window.h:56-65): channel = round_to_nearest_even(clamp(c * 255, 0, 255) - 0.5).
#pragma once
#include <cstdio>
#include <cstring>

#include <SDL2/SDL.h> // stub: keyboard state for Scene.cpp:165

#include "Math.h"
#include "Vector3.h"

struct Window {
	unsigned * frame_buffer;
	float    * frame_buffer_float; // extra: un-quantised linear colour, for error analysis only

	const int width;
	const int height;

	const int tile_width  = 32;
	const int tile_height = 32;

	const int tile_count_x;
	const int tile_count_y;

	bool is_closed = false;

	Window(int w, int h, const char * = nullptr) :
		width(w), height(h),
		tile_count_x((w + tile_width  - 1) / tile_width),
		tile_count_y((h + tile_height - 1) / tile_height)
	{
		frame_buffer       = new unsigned[size_t(w) * h]();
		frame_buffer_float = new float[size_t(w) * h * 3]();
	}
	~Window() { delete [] frame_buffer; delete [] frame_buffer_float; }

	void clear() { memset(frame_buffer, 0, size_t(width) * height * sizeof(unsigned)); }

	inline void plot(int x, int y, unsigned colour) const { frame_buffer[x + width * y] = colour; }

	inline void plot(int x, int y, const Vector3 & colour) const {
		float * f = frame_buffer_float + 3 * (size_t(x) + size_t(width) * y);
		f[0] = colour.x; f[1] = colour.y; f[2] = colour.z;

		int channel[3];
		for (int c = 0; c < 3; c++) {
			channel[c] = Util::float_to_int(Math::clamp(colour[c] * 255.0f, 0.0f, 255.0f) - 0.5f);
		}
		frame_buffer[x + width * y] = (channel[0] << 16) | (channel[1] << 8) | channel[2];
	}
};
