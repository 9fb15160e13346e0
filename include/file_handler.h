/*
 * file_handler.h -- image I/O helpers of the reference (file_handler.h:1-6,
 * file_handler.cpp:6-85) without OpenCV: images are binary PPM (P6) or PGM-free raw rasters.
 * tools/img2ppm.py converts the reference's PNG/JPG inputs.
 */
#pragma once
#include "header.h"

void loadImageIn(Mat &in_cam, vec3u *in_img, int height, int width);       /* BGR raster -> {s0=R,s1=G,s2=B} */
void loadImageOut(Mat &out_cam, cl_uchar3 *out_img, int height, int width);
/* one path per line; reads the first num_view images; returns true on ERROR (file_handler.cpp:30-57) */
bool read_image_array(vector<Mat> &img_array, std::string img_file_addr, int num_view);
void show_img(vec3u *res_img, int height, int width, int view_num);        /* writes a PPM, no window */
void imgSave(Mat img, int view_no);

bool read_ppm(const std::string &path, Mat &out);
bool write_ppm(const std::string &path, const Mat &img);
bool write_pgm(const std::string &path, const unsigned char *gray, int width, int height);
