/* Empty on purpose.  Used ONLY by `make -C sph_b200/host check-main` to syntax-check the reference's unmodified
 * main.cpp against the drop-in SphUtils.hpp (g++ -fsyntax-only: nothing is linked, nothing runs, no graphics API is
 * involved).  The reference's own vendored GLEW header, included first by main.cpp, already declares every GL type
 * and entry point; freeglut_std.h merely insists on including <GL/gl.h> and <GL/glu.h>, which this machine lacks. */
