// TEST INFRASTRUCTURE -- stand-in for <opencv/cv.hpp> so that the reference's CUDA/CUTracer.cu compiles unmodified.
// CUTracer.cu:383-396 uses OpenCV only to preview and save 8-bit step images inside RenderScene; none of it touches
// `hostcolor`.  Everything here is a no-op on a dummy image (no window, no files): headless box.
#pragma once
#define IPL_DEPTH_8U 8
struct IplImage { int width, height; };
struct CvSize { int width, height; };
struct CvScalar {
    double val[4];
    CvScalar(double a = 0, double b = 0, double c = 0, double d = 0) : val{a, b, c, d} {}
};
inline CvSize cvSize(int w, int h) { return CvSize{w, h}; }
inline IplImage* cvCreateImage(CvSize s, int, int) { return new IplImage{s.width, s.height}; }
inline void cvSet2D(IplImage*, int, int, CvScalar) {}
inline void cvShowImage(const char*, IplImage*) {}
inline int cvWaitKey(int) { return 0; }
inline int cvSaveImage(const char*, IplImage*) { return 1; }
inline void cvReleaseImage(IplImage** p) { delete *p; *p = nullptr; }
