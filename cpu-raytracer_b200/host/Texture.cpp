#include "Texture.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>

#include <zlib.h>

namespace b200rt {

static std::vector<uint8_t> read_file(const std::string & path) {
	FILE * f = fopen(path.c_str(), "rb");
	if (!f) throw std::runtime_error("cannot open image " + path);
	fseek(f, 0, SEEK_END);
	long size = ftell(f);
	rewind(f);
	std::vector<uint8_t> bytes(size > 0 ? size_t(size) : 0);
	size_t got = bytes.empty() ? 0 : fread(bytes.data(), 1, bytes.size(), f);
	fclose(f);
	if (got != bytes.size()) throw std::runtime_error("short read on image " + path);
	return bytes;
}

// ------------------------------------------------------------------------------- TGA ---
static Image decode_tga(const std::vector<uint8_t> & d, const std::string & path) {
	if (d.size() < 18) throw std::runtime_error("truncated TGA " + path);

	const int id_length = d[0], colour_map_type = d[1], image_type = d[2];
	const int width  = d[12] | (d[13] << 8);
	const int height = d[14] | (d[15] << 8);
	const int bpp    = d[16];
	const bool top_origin = (d[17] & 0x20) != 0;

	const bool rle = image_type == 10 || image_type == 11;
	const bool grey = image_type == 3 || image_type == 11;
	if (colour_map_type != 0 || !(image_type == 2 || image_type == 3 || rle)) throw std::runtime_error("unsupported TGA type in " + path);
	if (!((grey && bpp == 8) || (!grey && (bpp == 24 || bpp == 32)))) throw std::runtime_error("unsupported TGA depth in " + path);
	if (width <= 0 || height <= 0) throw std::runtime_error("empty TGA " + path);

	const int bytes_pp = bpp / 8;
	size_t pos = 18 + size_t(id_length);

	std::vector<uint8_t> raw(size_t(width) * height * bytes_pp);
	if (!rle) {
		if (pos + raw.size() > d.size()) throw std::runtime_error("truncated TGA " + path);
		memcpy(raw.data(), d.data() + pos, raw.size());
	} else {
		size_t out = 0;
		while (out < raw.size()) {
			if (pos >= d.size()) throw std::runtime_error("truncated RLE TGA " + path);
			int header = d[pos++];
			int count  = (header & 0x7f) + 1;
			if (header & 0x80) {
				if (pos + bytes_pp > d.size()) throw std::runtime_error("truncated RLE TGA " + path);
				for (int i = 0; i < count && out < raw.size(); i++, out += bytes_pp) memcpy(&raw[out], &d[pos], bytes_pp);
				pos += bytes_pp;
			} else {
				size_t n = std::min(size_t(count) * bytes_pp, raw.size() - out);
				if (pos + n > d.size()) throw std::runtime_error("truncated RLE TGA " + path);
				memcpy(&raw[out], &d[pos], n);
				pos += n; out += n;
			}
		}
	}

	Image image;
	image.width = width; image.height = height;
	image.rgba.resize(size_t(width) * height * 4);
	for (int y = 0; y < height; y++) {
		const int src_y = top_origin ? y : height - 1 - y;
		for (int x = 0; x < width; x++) {
			const uint8_t * s = &raw[(size_t(src_y) * width + x) * bytes_pp];
			uint8_t * o = &image.rgba[(size_t(y) * width + x) * 4];
			if (grey) { o[0] = o[1] = o[2] = s[0]; o[3] = 255; }
			else      { o[0] = s[2]; o[1] = s[1]; o[2] = s[0]; o[3] = bytes_pp == 4 ? s[3] : 255; } // stored BGR(A)
		}
	}
	return image;
}

// ------------------------------------------------------------------------------- PNG ---
static uint32_t be32(const uint8_t * p) { return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3]; }

static Image decode_png(const std::vector<uint8_t> & d, const std::string & path) {
	static const uint8_t signature[8] = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
	if (d.size() < 8 || memcmp(d.data(), signature, 8) != 0) throw std::runtime_error("not a PNG: " + path);

	int width = 0, height = 0, bit_depth = 0, colour_type = 0, interlace = 0;
	std::vector<uint8_t> idat, palette;

	size_t pos = 8;
	while (pos + 12 <= d.size()) {
		uint32_t length = be32(&d[pos]);
		const uint8_t * type = &d[pos + 4];
		const uint8_t * body = &d[pos + 8];
		if (pos + 12 + size_t(length) > d.size()) throw std::runtime_error("truncated PNG " + path);

		if (!memcmp(type, "IHDR", 4) && length >= 13) {
			width = int(be32(body)); height = int(be32(body + 4));
			bit_depth = body[8]; colour_type = body[9]; interlace = body[12];
		} else if (!memcmp(type, "PLTE", 4)) palette.assign(body, body + length);
		else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + length);
		else if (!memcmp(type, "IEND", 4)) break;

		pos += 12 + size_t(length);
	}

	int channels;
	switch (colour_type) {
		case 0: channels = 1; break; case 2: channels = 3; break; case 3: channels = 1; break;
		case 4: channels = 2; break; case 6: channels = 4; break;
		default: throw std::runtime_error("bad PNG colour type in " + path);
	}
	if (bit_depth != 8 || interlace != 0 || width <= 0 || height <= 0) throw std::runtime_error("unsupported PNG (need 8-bit, non-interlaced): " + path);

	const size_t stride = size_t(width) * channels;
	std::vector<uint8_t> raw((stride + 1) * height);

	uLongf raw_size = uLongf(raw.size());
	if (uncompress(raw.data(), &raw_size, idat.data(), uLong(idat.size())) != Z_OK || raw_size != raw.size()) throw std::runtime_error("PNG inflate failed: " + path);

	// Undo the per-scanline filters (PNG spec 9.2)
	std::vector<uint8_t> pixels(stride * height);
	for (int y = 0; y < height; y++) {
		const uint8_t filter = raw[(stride + 1) * y];
		const uint8_t * in   = &raw[(stride + 1) * y + 1];
		uint8_t * out        = &pixels[stride * y];
		const uint8_t * up   = y ? &pixels[stride * (y - 1)] : nullptr;

		for (size_t i = 0; i < stride; i++) {
			int a = i >= size_t(channels) ? out[i - channels] : 0;
			int b = up ? up[i] : 0;
			int c = (up && i >= size_t(channels)) ? up[i - channels] : 0;
			int predicted;
			switch (filter) {
				case 0: predicted = 0; break;
				case 1: predicted = a; break;
				case 2: predicted = b; break;
				case 3: predicted = (a + b) >> 1; break;
				case 4: { int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
				          predicted = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); } break;
				default: throw std::runtime_error("bad PNG filter in " + path);
			}
			out[i] = uint8_t(in[i] + predicted);
		}
	}

	Image image;
	image.width = width; image.height = height;
	image.rgba.resize(size_t(width) * height * 4);
	for (size_t p = 0; p < size_t(width) * height; p++) {
		const uint8_t * s = &pixels[p * channels];
		uint8_t * o = &image.rgba[p * 4];
		switch (colour_type) {
			case 0: o[0] = o[1] = o[2] = s[0]; o[3] = 255; break;
			case 2: o[0] = s[0]; o[1] = s[1]; o[2] = s[2]; o[3] = 255; break;
			case 3: if (size_t(s[0]) * 3 + 2 >= palette.size()) throw std::runtime_error("PNG palette index out of range: " + path);
			        o[0] = palette[s[0] * 3]; o[1] = palette[s[0] * 3 + 1]; o[2] = palette[s[0] * 3 + 2]; o[3] = 255; break;
			case 4: o[0] = o[1] = o[2] = s[0]; o[3] = s[1]; break;
			case 6: memcpy(o, s, 4); break;
		}
	}
	return image;
}

Image load_image(const std::string & path) {
	std::vector<uint8_t> bytes = read_file(path);
	if (bytes.size() >= 8 && bytes[0] == 0x89 && bytes[1] == 'P') return decode_png(bytes, path);
	return decode_tga(bytes, path);
}

void save_tga(const std::string & path, int width, int height, const uint8_t * rgb) {
	FILE * f = fopen(path.c_str(), "wb");
	if (!f) throw std::runtime_error("cannot write " + path);
	uint8_t header[18] = { 0 };
	header[2] = 2;
	header[12] = uint8_t(width & 0xff);  header[13] = uint8_t(width >> 8);
	header[14] = uint8_t(height & 0xff); header[15] = uint8_t(height >> 8);
	header[16] = 24; header[17] = 0x20; // top-left origin
	fwrite(header, 1, 18, f);
	std::vector<uint8_t> row(size_t(width) * 3);
	for (int y = 0; y < height; y++) {
		for (int x = 0; x < width; x++) {
			const uint8_t * s = rgb + (size_t(y) * width + x) * 3;
			row[x * 3] = s[2]; row[x * 3 + 1] = s[1]; row[x * 3 + 2] = s[0];
		}
		fwrite(row.data(), 1, row.size(), f);
	}
	fclose(f);
}

// ------------------------------------------------------------------------- mip pyramid ---
float gamma_to_linear(float x) {
	if (x <= 0.0f) return 0.0f;
	if (x >= 1.0f) return 1.0f;
	if (x < 0.04045f) return x / 12.92f;
	return powf((x + 0.055f) / 1.055f, 2.4f);
}

static bool is_power_of_two(int n) { return n != 0 && (n & (n - 1)) == 0; }

int TexturePool::load(const std::string & path) {
	auto it = cache.find(path);
	if (it != cache.end()) return it->second;
	int index = add(load_image(path));
	cache[path] = index;
	return index;
}

TexturePool::TexturePool() {
	for (int i = 0; i < 256; i++) gamma_lut[i] = gamma_to_linear(float(i) * 0.00392156862f);
}

int TexturePool::add(const Image & image) {
	const int w = image.width, h = image.height;

	b200rt_texture texture;
	memset(&texture, 0, sizeof(texture));
	texture.texel_offset = int64_t(texels.size() / 3);
	texture.width  = w;
	texture.height = h;
	// Only power-of-two textures get a pyramid (Texture.cpp:50)
	texture.mipmapped  = is_power_of_two(w) && is_power_of_two(h);
	texture.mip_levels = texture.mipmapped ? 1 + int(log2f(float(std::min(w, h)))) : 1; // Texture.cpp:77
	if (texture.mip_levels > B200RT_MAX_MIP_LEVELS) throw std::runtime_error("texture too large");

	// Level 0: byte/255 (as a multiply by 0.00392156862f, Texture.cpp:13-20) then sRGB -> linear.
	// A 256-entry table gives the same floats as converting texel by texel.
	const float * lut = gamma_lut;

	const size_t base = texels.size();
	texels.resize(base + size_t(w) * h * 3);
	const size_t base8 = rgb8.size();
	rgb8.resize(base8 + size_t(w) * h * 3);
	for (size_t p = 0; p < size_t(w) * h; p++) {
		for (int c = 0; c < 3; c++) {
			rgb8[base8 + p * 3 + c] = image.rgba[p * 4 + c];
			texels[base + p * 3 + c] = lut[image.rgba[p * 4 + c]];
		}
	}

	if (texture.mipmapped) {
		// Each level is the 2x2 box filter of the previous one; the chain stops when either side
		// reaches zero (Texture.cpp:93), so a 256x1024 texture ends at 1x4.
		int prev_w = w, prev_h = h;
		size_t prev_offset = 0, offset = size_t(w) * h;
		int level = 1;
		for (int lw = w >> 1, lh = h >> 1; lw >= 1 && lh >= 1; lw >>= 1, lh >>= 1) {
			texels.resize(base + (offset + size_t(lw) * lh) * 3);
			float * data = texels.data() + base;
			for (int j = 0; j < lh; j++) {
				for (int i = 0; i < lw; i++) {
					const float * c0 = &data[(prev_offset + size_t(2 * i)     + size_t(2 * j)     * prev_w) * 3];
					const float * c1 = &data[(prev_offset + size_t(2 * i + 1) + size_t(2 * j)     * prev_w) * 3];
					const float * c2 = &data[(prev_offset + size_t(2 * i)     + size_t(2 * j + 1) * prev_w) * 3];
					const float * c3 = &data[(prev_offset + size_t(2 * i + 1) + size_t(2 * j + 1) * prev_w) * 3];
					float * out = &data[(offset + size_t(i) + size_t(j) * lw) * 3];
					for (int c = 0; c < 3; c++) out[c] = (((c0[c] + c1[c]) + c2[c]) + c3[c]) * 0.25f;
				}
			}
			if (level < B200RT_MAX_MIP_LEVELS) texture.mip_offsets[level] = int32_t(offset);
			level++;
			prev_offset = offset; offset += size_t(lw) * lh;
			prev_w = lw; prev_h = lh;
		}
		(void)prev_h;
	}

	textures.push_back(texture);
	return int(textures.size()) - 1;
}

void ewa_weight_table(float table[128]) {
	const float alpha = 2.0f;
	const float denom = 1.0f / 127.0f;
	const float exp_neg_alpha = expf(-alpha);
	for (int i = 0; i < 128; i++) {
		float r2 = float(i) * denom;
		table[i] = expf(-alpha * r2) - exp_neg_alpha;
	}
}

} // namespace b200rt
