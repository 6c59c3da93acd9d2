// TEST INFRASTRUCTURE — stub of the one SDL2 call and the scancodes the reference's
// Camera.cpp / Scene.cpp touch (Camera.cpp:26-39, Scene.cpp:165). No window, no input:
// the keyboard is always "nothing pressed", so Camera::update only rotates the view
// pyramid (Camera.cpp:44-47).
#pragma once
enum {
	SDL_SCANCODE_A = 4, SDL_SCANCODE_D = 7, SDL_SCANCODE_F = 9, SDL_SCANCODE_S = 22, SDL_SCANCODE_W = 26,
	SDL_SCANCODE_SPACE = 44, SDL_SCANCODE_RIGHT = 79, SDL_SCANCODE_LEFT = 80, SDL_SCANCODE_DOWN = 81,
	SDL_SCANCODE_UP = 82, SDL_SCANCODE_LSHIFT = 225, SDL_NUM_SCANCODES = 512
};
inline const unsigned char * SDL_GetKeyboardState(int *) {
	static unsigned char keys[SDL_NUM_SCANCODES] = { 0 };
	return keys;
}
