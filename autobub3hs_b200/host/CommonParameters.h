// reference: common/CommonParameters.h:5
#ifndef ABUB_COMMONPARAMETERS_H
#define ABUB_COMMONPARAMETERS_H
#define NumFramesBubbleTrack 10
#endif
