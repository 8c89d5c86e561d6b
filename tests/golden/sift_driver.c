/* sift_driver.c -- data generator for the C2 fixture (castle image pair), NOT part of
 * the product: runs vlfeat's SIFT (vendored by the reference under /root/reference/vlfeat,
 * compiled where it lies by make_castle.py) over a grayscale float image and returns one
 * row per (keypoint, orientation): x, y, sigma, angle, then the 128 descriptor bins as
 * (uint8)(512 * bin) -- the row format of the reference's sift_filter
 * (src/Sift.h:49-137 with its defaults O=-1, S=3, omin=-1, :37-42).
 * Validated against data/sift-test/sur-ogre.sift the way test/test_feature.py:32-47 does. */
#include <stdlib.h>
#include <string.h>
#include <vl/sift.h>

/* returns the number of rows; *out is malloc'ed (rows x 132 floats), caller frees with sift_free */
int sift_rows(const float *image, int width, int height, float **out) {
  int cap = 4096, n = 0;
  float *rows = (float *)malloc((size_t)cap * 132 * sizeof(float));
  VlSiftFilt *f = vl_sift_new(width, height, -1, 3, -1);
  if (!f || !rows) return -1;
  int first = 1;
  for (;;) {
    int err = first ? vl_sift_process_first_octave(f, image) : vl_sift_process_next_octave(f);
    first = 0;
    if (err) break;
    vl_sift_detect(f);
    const VlSiftKeypoint *kp = vl_sift_get_keypoints(f);
    const int nkp = vl_sift_get_nkeypoints(f);
    for (int i = 0; i < nkp; ++i) {
      double ang[4];
      const int na = vl_sift_calc_keypoint_orientations(f, ang, kp + i);
      for (int a = 0; a < na; ++a) {
        vl_sift_pix d[128];
        vl_sift_calc_keypoint_descriptor(f, d, kp + i, ang[a]);
        if (n == cap) {
          cap *= 2;
          rows = (float *)realloc(rows, (size_t)cap * 132 * sizeof(float));
          if (!rows) return -1;
        }
        float *r = rows + (size_t)n * 132;
        r[0] = kp[i].x; r[1] = kp[i].y; r[2] = kp[i].sigma; r[3] = (float)ang[a];
        for (int b = 0; b < 128; ++b) r[4 + b] = (float)(vl_uint8)(512.0 * d[b]);
        ++n;
      }
    }
  }
  vl_sift_delete(f);
  *out = rows;
  return n;
}

void sift_free(float *p) { free(p); }
