// Host-side driver for tests/test_glm_standin.py: evaluates the glm stand-in (oracle/glm_standin/glm/glm.hpp) on the
// matrices / vectors read from stdin and prints the results, so that numpy can pin its conventions.
#include <cstdio>
#include <glm/glm.hpp>

int main() {
    int n = 0;
    if (scanf("%d", &n) != 1) return 1;
    for (int t = 0; t < n; ++t) {
        float a[9], b[9], v[3], w[3];
        for (float& x : a) if (scanf("%f", &x) != 1) return 1;
        for (float& x : b) if (scanf("%f", &x) != 1) return 1;
        for (float& x : v) if (scanf("%f", &x) != 1) return 1;
        for (float& x : w) if (scanf("%f", &x) != 1) return 1;
        // the 9-scalar constructor fills COLUMNS (glm is column major): a[0..2] is column 0
        const glm::mat3 A(a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8]);
        const glm::mat3 B(b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7], b[8]);
        const glm::vec3 V(v[0], v[1], v[2]), W(w[0], w[1], w[2]);
        const glm::mat3 AB = A * B, At = glm::transpose(A), sA = 2.5f * A;
        const glm::vec3 AV = A * V, mx = glm::max(V, 0.0f);
        for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) printf("%.9g ", AB[j][i]);      // column by column
        for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) printf("%.9g ", At[j][i]);
        for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) printf("%.9g ", sA[j][i]);
        printf("%.9g %.9g %.9g ", AV.x, AV.y, AV.z);
        printf("%.9g %.9g ", glm::dot(V, W), glm::length(V));
        printf("%.9g %.9g %.9g\n", mx.x, mx.y, mx.z);
    }
    return 0;
}
