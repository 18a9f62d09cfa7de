/*
 * node_io_oracle.c -- TEST INFRASTRUCTURE (see amcl_oracle.c for the rules).
 * CPU restatement of the steps either side of the AMCL filter that SURVEY.md
 * 8(f) ranks next: map ingest (N4), the published particle cloud (N2) and
 * costmap inflation (N3).  Paths are under /root/reference.
 *
 * Parity status
 *   - orc_image_to_grid: PINNED by the golden vectors of the reference's own
 *     unit tests (map_server/test/test_constants.cpp: testmap.png / .bmp /
 *     testmap2.png contents; map_server/test/utest.cpp:121-152: value counts
 *     of spectrum.png in the three modes) -- tests/golden/map_server_golden.npz,
 *     written by tests/golden/make_map_golden.py.  The loader itself cannot be
 *     compiled here (SDL_image, nav_msgs).
 *   - orc_convert_map: trivially restated from amcl_node.cpp:1002-1009 (the node
 *     cannot be compiled here: roscpp).  Parity unpinned by execution.
 *   - orc_pack_cloud: amcl_node.cpp:1349-1362 with tf2::Quaternion::setRPY
 *     restated from its published formula (ros/geometry2, tf2/LinearMath/
 *     Quaternion.h, the 0.6/0.7 series ROS melodic/noetic ship; tf2 is a
 *     package dependency of amcl, not vendored in /root/reference).  Parity
 *     unpinned by execution.
 *   - orc_inflate: costmap_2d/plugins/inflation_layer.cpp:176-306; see there.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { ORC_TRINARY = 0, ORC_SCALE = 1, ORC_RAW = 2 }; /* map_server/image_loader.h:46 */

/* map_server/src/image_loader.cpp:104-164.  pixels: rowstride bytes per image
 * row, n_channels bytes per pixel, top row first; grid: width*height values,
 * bottom row first. */
void orc_image_to_grid(const unsigned char *pixels, int width, int height, int rowstride,
                       int n_channels, int has_alpha_mask, int negate, double occ_th,
                       double free_th, int mode, signed char *grid)
{
  int avg_channels;
  if (mode == ORC_TRINARY || !has_alpha_mask) /* :106-109 */
    avg_channels = n_channels;
  else
    avg_channels = n_channels - 1;
  for (int j = 0; j < height; j++)
    for (int i = 0; i < width; i++)
    {
      const unsigned char *p = pixels + (size_t)j * rowstride + (size_t)i * n_channels;
      unsigned char value;
      int color_sum = 0;
      for (int k = 0; k < avg_channels; k++) color_sum += p[k];
      double color_avg = color_sum / (double)avg_channels;
      int alpha = n_channels == 1 ? 1 : p[n_channels - 1];
      if (negate) color_avg = 255 - color_avg;
      if (mode == ORC_RAW)
      {
        value = color_avg;
        grid[(size_t)width * (height - j - 1) + i] = value;
        continue;
      }
      double occ = (255 - color_avg) / 255.0;
      if (occ > occ_th)
        value = +100;
      else if (occ < free_th)
        value = 0;
      else if (mode == ORC_TRINARY || alpha < 1.0)
        value = -1;
      else
      {
        double ratio = (occ - free_th) / (occ_th - free_th);
        value = 1 + 98 * ratio;
      }
      grid[(size_t)width * (height - j - 1) + i] = value;
    }
}

/* amcl/src/amcl_node.cpp:1002-1009 */
void orc_convert_map(const signed char *grid, size_t n, int8_t *occ_state)
{
  for (size_t i = 0; i < n; i++)
  {
    if (grid[i] == 0)
      occ_state[i] = -1;
    else if (grid[i] == 100)
      occ_state[i] = +1;
    else
      occ_state[i] = 0;
  }
}

/* amcl/src/amcl_node.cpp:1349-1362; samples are rows (x, y, theta, w).
 * out: 7 doubles per pose (position xyz, orientation xyzw). */
int orc_pack_cloud(const double *samples, int n, int stride, int max_poses, double *out)
{
  int count = 0;
  for (int i = 0; i < n && count < max_poses; i += stride, count++)
  {
    const double roll = 0, pitch = 0, yaw = samples[4 * (size_t)i + 2];
    /* tf2::Quaternion::setRPY */
    double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
    double cosYaw = cos(halfYaw), sinYaw = sin(halfYaw);
    double cosPitch = cos(halfPitch), sinPitch = sin(halfPitch);
    double cosRoll = cos(halfRoll), sinRoll = sin(halfRoll);
    double *o = out + 7 * (size_t)count;
    o[0] = samples[4 * (size_t)i + 0];
    o[1] = samples[4 * (size_t)i + 1];
    o[2] = 0;
    o[3] = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
    o[4] = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
    o[5] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
    o[6] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  }
  return count;
}
