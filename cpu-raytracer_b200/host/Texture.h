// Host side of the texture path: decode an image file, convert sRGB bytes to linear float
// RGB and build the box-filtered mip pyramid the sampling kernels read
// (Texture::load, Texture.cpp:30-129). All textures of a scene live in one texel pool that
// is uploaded to HBM once; a b200rt_texture describes one pyramid inside the pool.
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/b200rt.h"

namespace b200rt {

// 8-bit RGBA image, row 0 = top (what stb_image hands the reference, Texture.cpp:39).
struct Image {
	int width = 0, height = 0;
	std::vector<uint8_t> rgba;
};

// Decoders for the formats the reference's assets use: Truevision TGA (24/32-bit, raw or RLE,
// either origin) and PNG (8-bit grey / grey+alpha / RGB / RGBA / palette, non-interlaced).
// Throws std::runtime_error on anything else.
Image load_image(const std::string & path);

// Minimal writer (uncompressed 24-bit, top-left origin) for procedurally generated assets.
void save_tga(const std::string & path, int width, int height, const uint8_t * rgb);

struct TexturePool {
	std::vector<b200rt_texture> textures;
	std::vector<float>          texels; // float[n][3]: every pyramid, built on the host
	std::vector<uint8_t>        rgb8;   // level-0 images only, 3 bytes per texel, texture after texture: input of the GPU ingest
	float                       gamma_lut[256]; // Math::gamma_to_linear(i * 0.00392156862f)

	TexturePool();

	// Returns the texture index; a path is loaded once (texture_cache, Texture.cpp:10,31-34).
	int load(const std::string & path);
	int add(const Image & image);

private:
	std::unordered_map<std::string, int> cache;
};

float gamma_to_linear(float x); // Math.h:67-77

// 128-entry EWA Gaussian weight table, alpha = 2 (Texture::init, Texture.h:53-62).
void ewa_weight_table(float table[128]);

} // namespace b200rt
