// reference: bubble/bubble.cpp.  Float/double mix kept as written there, because dZdt/dRdt are
// printed to two decimals (and as nan/-nan for untracked bubbles: 0/0).
#include "bubble.hpp"

#include <math.h>

#include <iostream>

bubble::bubble(BubbleImageFrame g) {                       // bubble.cpp:34-48
    KnownDescriptors.push_back(g);
    last_x = g.MassCentres.x;
    last_y = g.MassCentres.y;
    GenesisPositionCentroid = g.MassCentres;
    GenesisPosition = g.newPosition;
    lockThisIteration = true;
}

bubble::~bubble() {}

void bubble::operator<<(BubbleImageFrame f) {              // bubble.cpp:54-72: appends only when unlocked
    if (lockThisIteration) return;
    KnownDescriptors.push_back(f);
    dz.push_back(last_x - f.newPosition.x);
    last_x = f.MassCentres.x;
    last_y = f.MassCentres.y;
    lockThisIteration = true;
}

void bubble::printAllXY(void) {
    for (const BubbleImageFrame& d : KnownDescriptors) std::cout << "X: " << d.MassCentres.x << " ";
    std::cout << "\n";
    for (const BubbleImageFrame& d : KnownDescriptors) std::cout << "Y: " << d.MassCentres.y << " ";
    std::cout << "\n";
}

bool bubble::isNewPositionProbable(int& x, int& y) {       // bubble.cpp:88-95 (unused by the live path)
    return fabs(last_y - y) <= 4 && (last_x - x) < 5;
}

float bubble::dZdT(void) {                                 // bubble.cpp:101-108
    const int numFrames = (int)KnownDescriptors.size();
    volatile float total_z = (float)(KnownDescriptors[0].newPosition.x - KnownDescriptors[numFrames - 1].newPosition.x);
    volatile double den = (double)(float)numFrames - 1.0;  // volatile: 0/0 must be evaluated by the FPU (-nan), not folded
    return (float)((double)total_z / den);
}

float bubble::dRdT(void) {                                 // bubble.cpp:110-118
    const int numFrames = (int)KnownDescriptors.size();
    const float dx = (float)(KnownDescriptors[0].newPosition.width - KnownDescriptors[numFrames - 1].newPosition.width);
    const float dy = (float)(KnownDescriptors[0].newPosition.height - KnownDescriptors[numFrames - 1].newPosition.height);
    volatile float dr = sqrtf(dx * dx + dy * dy);
    volatile double den = (double)(float)numFrames - 1.0;
    return (float)((double)dr / den);
}
