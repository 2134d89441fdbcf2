/* oracle/ref_stb/stb_decode.cpp -- TEST INFRASTRUCTURE.  The reference's own image decoder (the stb_image.h it vendors,
 * compiled where it lies: -DREF_STB='"/root/reference/include/stb/stb_image.h"') behind the two steps of
 * loadTexture (/root/reference/src/Model.cpp:99-118): stbi_load(file, &w, &h, &comp, STBI_rgb_alpha), then the mirror
 * along y.  Prints "w h\n" and the w*h*4 RGBA8 bytes (first row = the image's BOTTOM row) to stdout.
 * tests/golden/make_golden_textures.py turns its output into the fixtures host/image_io.cpp is pinned against. */
#define STB_IMAGE_IMPLEMENTATION
#include REF_STB
#include <stdio.h>
#include <string.h>

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    int w = 0, h = 0, comp = 0;
    unsigned char* px = stbi_load(argv[1], &w, &h, &comp, STBI_rgb_alpha);
    if (!px) {
        fprintf(stderr, "stbi_load failed: %s\n", stbi_failure_reason());
        return 1;
    }
    for (int y = 0; y < h / 2; y++) { /* Model.cpp:108-116 */
        unsigned char* a = px + (size_t)4 * w * y;
        unsigned char* b = px + (size_t)4 * w * (h - 1 - y);
        for (int i = 0; i < 4 * w; i++) {
            unsigned char t = a[i];
            a[i] = b[i];
            b[i] = t;
        }
    }
    printf("%d %d\n", w, h);
    fwrite(px, 1, (size_t)4 * w * h, stdout);
    return 0;
}
