// intentionally empty: the reference includes stb_image.h but never calls it on this path
