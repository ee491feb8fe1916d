// oracle/matchshim: empty stand-in for <boost/serialization/serialization.hpp> (test infrastructure); names come from facade.h
